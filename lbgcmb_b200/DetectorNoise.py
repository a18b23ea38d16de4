"""Instrumental noise spectrum; mirrors reference cmb/DetectorNoise.py:8-32.

Elementwise on <= 1e4 table entries when called from Python (the N0 kernel evaluates the same
expression per integrand node on the device, csrc/n0.cu)."""
import numpy as np

from lbgcmb_b200.params import get_params

params = get_params()


def DetectorNoise(ell, thetab=1., DeltaT=1., type='TT'):
    DeltaT = DeltaT * 10. ** -6.           # micro K -> K (the reference rebinds its local copy the same way)
    arcmin2radians = np.pi / (60. * 180.)
    DeltaT = DeltaT * arcmin2radians       # K-arcmin -> K-radian
    thetab = thetab * arcmin2radians
    DeltaP = np.sqrt(2.) * DeltaT          # fully polarised detectors
    if type == 'TT':
        return (DeltaT / params['Tcmb']) ** 2. * np.exp(ell * (ell + 1.) * thetab ** 2. / (8. * np.log(2.)))
    elif type in ['EE', 'BB', 'EB']:
        return (DeltaP / params['Tcmb']) ** 2. * np.exp(ell * (ell + 1.) * thetab ** 2. / (8. * np.log(2.)))
    elif type == 'TE':
        return 0.0
    else:
        raise ValueError("Noise estimate is not available for %s combination." % type)
