"""P(k,z) provider glue; mirrors reference pmh.py:8-47 (`get_PkInterps`, `Pmm`, `Pmh`, `Phh`, bias helpers).

The reference takes P(k,z) from `camb` interpolators (prep_camb.py:55-68); camb is absent and unpinned,
so the provider here is the reference's *other*, fully in-repo implementation: linear table x D(z)^2 -> HALOFIT for
z <= 4 (modi/cosmology.py:158-182, modi/halofit.py) -> np.interp in k, zero outside the table.  `PkTable` plays the role
of the `CAMB` object (same two getter names); the dict values are table-backed callables `(z, k[]) -> P[]` evaluated on
the GPU.  A foreign callable is rejected by the Limber mirrors with TypeError: there is no CPU fallback.
"""
import numpy as np

from lbgcmb_b200.growth_rate import growth_factor
from lbgcmb_b200.modi.cosmology import Cosmology
from lbgcmb_b200.params import get_params


def unit_bz(z):
    return np.ones_like(z)


def linz_bz(z):
    return (1. + z)


def invG_bz(z):
    return 1. / growth_factor(z)


class PkCallable(object):
    """`Pk_interps[...]`: called as (z: float, k: ndarray) -> ndarray, like np.vectorize(PKInterpolator.P)."""

    def __init__(self, engine, nonlinear):
        self.engine = engine
        self.nonlinear = nonlinear

    def __call__(self, z, k):
        ka = np.asarray(k, dtype=np.float64)
        out = self.engine.pk_eval(float(z), np.atleast_1d(ka), nonlinear=self.nonlinear, oob=0)
        return out.reshape(ka.shape) if ka.ndim else float(out[0])


class PkTable(object):
    """Stands where `prep_camb.CAMB()` stands in the reference drivers (Ckg.py:98-104): holds the linear table and
    hands out the two (z, k) interpolators."""

    def __init__(self, k, plin_z0, params=None, device=None):
        params = get_params() if params is None else params
        self.engine = Cosmology(M=params['om_m'], H0=100., B=params['om_b'], sig8=params['sig_8'],
                                klin=np.asarray(k, dtype=np.float64), plin=np.asarray(plin_z0, dtype=np.float64), device=device)

    def get_lpower_interpolatekz(self, zscatter=None):
        return PkCallable(self.engine, False)

    def get_nlpower_interpolatekz(self, zscatter=None):
        return PkCallable(self.engine, True)


def get_PkInterps(cambx):
    Pk_interps = {}
    Pk_interps['camb_lin'] = cambx.get_lpower_interpolatekz()
    Pk_interps['camb_nlin'] = cambx.get_nlpower_interpolatekz()
    return Pk_interps


def engine_of(Pk_interps, type='nlinear'):
    """(engine, nonlinear flag) behind a Pk_interps dict, or TypeError for foreign callables."""
    if type == 'linear':
        f = Pk_interps['camb_lin']
    elif type == 'nlinear':
        f = Pk_interps['camb_nlin']
    else:
        raise ValueError('Type %s not available for Pmm.' % type)
    if not isinstance(f, PkCallable):
        raise TypeError('Pk_interps entries must be table-backed PkCallable objects (pmh.PkTable); '
                        'arbitrary Python callables cannot run on the GPU and there is no CPU fallback')
    return f.engine, f.nonlinear


def Pmm(Pk_interps, k, z, type='nlinear'):
    if type == 'linear':
        result = Pk_interps['camb_lin'](z, k)
    elif type == 'nlinear':
        result = Pk_interps['camb_nlin'](z, k)
    else:
        raise ValueError('Type %s not available for Pmm.' % type)
    return result


def Pmh(Pk_interps, bz, k, z):
    return bz(z) * Pk_interps['camb_nlin'](z, k)


def Phh(Pk_interps, bz, k, z, bz2=None):
    if bz2 is None:
        return Pk_interps['camb_nlin'](z, k) * bz(z) ** 2.
    else:
        return Pk_interps['camb_nlin'](z, k) * bz(z) * bz2(z)
