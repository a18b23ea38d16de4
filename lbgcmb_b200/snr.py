"""Cumulative signal-to-noise; mirrors reference snr.py:21-28 (`snr`, `zeros`).  The masked sum over multipoles runs on
the device (csrc/fisher.cu snr_kernel) through lbg_snr."""
import numpy as np

from lbgcmb_b200 import _ffi


def snr(Llls, Cls, Vars, nmodes, lmax=1.e5, device=None):
    ctx = _ffi.context(device)
    L, C, V, N = (_ffi.f64(x) for x in (Llls, Cls, Vars, nmodes))
    if not (L.shape == C.shape == V.shape == N.shape) or L.ndim != 1:
        raise ValueError('Llls, Cls, Vars and nmodes must be 1-d arrays of one length')
    out = np.empty(1)
    _ffi.check(_ffi.lib().lbg_snr(ctx.handle, L.size, _ffi._dp(L), _ffi._dp(C), _ffi._dp(V), _ffi._dp(N), float(lmax), _ffi._dp(out)))
    return np.float64(out[0])


def zeros(arg):
    return np.zeros_like(arg)
