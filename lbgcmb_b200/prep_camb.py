"""CMB C_l containers; mirrors the parts of reference prep_camb.py the hot path touches.

prep_camb.py:70-100 (`CAMB.get_Cls`) returns two dicts {'TT','EE','BB','TE','kk'} -> cubic interp1d on
integer l, zero outside the table.  CAMB itself is absent, so here the dicts are built from
in-memory tables (`cl_interps_from_tables`); each value is a `ClInterp`, a table-backed callable whose
evaluation is the device spline kernel (csrc/ctx.cu lbg_eval_spline) -- same not-a-knot spline as interp1d.
`Clxy` mirrors prep_camb.py:128-137.
"""
import numpy as np

from lbgcmb_b200 import _ffi

SPECTRA = ('TT', 'EE', 'BB', 'TE')   # prep_camb.py:84 column order


class ClInterp(object):
    """interp1d(ell, cl, kind='cubic', bounds_error=False, fill_value=0.0) with ell = 0..n-1, on the GPU."""

    def __init__(self, ell, cl):
        ell = np.asarray(ell, dtype=np.float64)
        if ell.ndim != 1 or ell.size < 4 or not np.array_equal(ell, np.arange(ell.size)):
            raise ValueError('ClInterp needs integer knots ell = 0..n-1 (prep_camb.py:80)')
        self.x = ell
        self.y = _ffi.f64(cl)
        if self.y.shape != ell.shape:
            raise ValueError('ell and cl differ in shape')

    def __call__(self, x):
        xa = _ffi.f64(np.atleast_1d(x))
        out = np.empty(xa.size)
        ctx = _ffi.context()
        _ffi.check(_ffi.lib().lbg_eval_spline(ctx.handle, self.y.size, _ffi._dp(self.y), xa.size,
                                              _ffi._dp(xa.ravel()), _ffi._dp(out)))
        out = out.reshape(xa.shape)
        return out if np.ndim(x) else out[0]


def cl_interps_from_tables(ell, lensed, unlensed, kk=None):
    """(lensCl_interps, nolensCl_interps) as CAMB.get_Cls() returns them (prep_camb.py:86-100).

    lensed / unlensed: [4, n] arrays in TT,EE,BB,TE order (raw, dimensionless C_l), or dicts by name."""
    def mk(tab):
        if isinstance(tab, dict):
            return {x: ClInterp(ell, tab[x]) for x in SPECTRA}
        return {x: ClInterp(ell, tab[i]) for i, x in enumerate(SPECTRA)}
    lens, nolens = mk(lensed), mk(unlensed)
    if kk is not None:
        lens['kk'] = nolens['kk'] = ClInterp(ell, kk)
    return lens, nolens


def table_of(interp, n_default=None):
    """Knot values of a C_l callable on l = 0..n-1: its own table if it has one, else sampled."""
    x = getattr(interp, 'x', None)
    y = getattr(interp, 'y', None)
    if x is not None and y is not None:
        if isinstance(interp, ClInterp):                       # integer knots 0..n-1 checked at construction
            return interp.y
        x = np.asarray(x, dtype=np.float64)
        if x.ndim == 1 and np.array_equal(x, np.arange(x.size)):
            return _ffi.f64(y)
    if n_default is None:
        raise TypeError('C_l callable without an integer-knot table; pass table-backed interpolators')
    return _ffi.f64(interp(np.arange(n_default, dtype=np.float64)))


def Clxy(Cl_interps, ell, alpha='TT', thetab=1., DeltaT=1.):
    """C_l + detector noise, prep_camb.py:128-137."""
    from lbgcmb_b200.DetectorNoise import DetectorNoise
    result = Cl_interps[alpha](ell)
    result = result + DetectorNoise(ell, thetab, DeltaT, type=alpha)
    return result
