"""lbgcmb_b200 -- B200-native (sm_100a CUDA behind a C ABI) implementation of the LBGCMB forecast hot path.

The modules mirror the reference's Python call surface:
  Nkk.Nkk, DetectorNoise.DetectorNoise, prep_camb.Clxy, Cgg.Cgg/Ngg/var_Cgg, Ckg.Ckg/var_Ckg,
  lensing.lensing_kernel/Ckk/var_Ckk, pmh.Pmm/Pmh/Phh, snr.snr, prep_Llls.prep_Llls, ...
  modi.nkk, modi.halofit, modi.cosmology, modi.lensing
There is NO CPU fallback: every hot function goes through `_ffi.lib()` (liblbgcmb.so) and raises
if the library or a CUDA device is missing.
"""
__version__ = '0.1.0'
