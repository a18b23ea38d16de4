"""Linear b(z) of the Schmittfull & Seljak (2017) tracers; mirrors reference schmittfull_bz.py:10-111 (host-side)."""
import os
from collections import OrderedDict

import numpy as np
from scipy.interpolate import interp1d

from lbgcmb_b200.growth_rate import growth_factor


def sdss(zee, force=True):
    zee = np.asarray(zee)
    return 1.0 + (zee - 0.1)


def _growth_bias(amp, zmax, name):
    def bz(zee, force=False):
        zee = np.asarray(zee)
        aa = 1. / (1. + zee)
        if force:
            return amp / growth_factor(aa)
        if np.any(zee > zmax):
            raise ValueError(name)
        return amp / growth_factor(aa)
    return bz


boss_lrg = _growth_bias(1.70, 1.0, '\n\nCalling b(z) outside range of BOSS LRGs (z < 1.0).\n\n')
desi_bgs = _growth_bias(1.34, 0.5, '\n\nCalling b(z) outside range of BGS (z < 0.5).\n\n')
desi_lrg = _growth_bias(1.70, 1.3, '\n\nCalling b(z) outside range of LRGs (z < 1.3).\n\n')
desi_elg = _growth_bias(0.84, 1.8, '\n\nCalling b(z) outside range of ELGs (z < 1.8).\n\n')
desi_qso = _growth_bias(1.20, 5.0, '\n\nCalling b(z) outside range of QSOs (z < 5.0).\n\n')


def desi_eff(zee):
    root = os.environ['LBGCMB']
    data = np.loadtxt(root + '/cadence/dat/effective_desi.dat')
    ibz = interp1d(data[:, 0], data[:, 2], kind='linear', bounds_error=False, fill_value=0.0)
    return ibz(np.asarray(zee))


def get_allbz():
    """Survey name (dat/ss17_tracers.dat) -> b(z); schmittfull_bz.py:87-111."""
    result = OrderedDict()
    result['SDSS1'] = lambda z: sdss(z, force=True)
    result['SDSS2'] = lambda z: sdss(z, force=True)
    result['DESIBGS'] = lambda z: desi_bgs(z, force=True)
    result['BOSSLRGs'] = lambda z: boss_lrg(z, force=True)
    result['DESILRGs'] = lambda z: desi_lrg(z, force=True)
    result['DESIELGs1'] = lambda z: desi_elg(z, force=True)
    result['DESIELGs2'] = lambda z: desi_elg(z, force=True)
    result['DESIQSOs'] = lambda z: desi_qso(z, force=True)
    return result
