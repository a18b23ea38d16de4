"""Limber C_kg and its variance; mirrors reference Ckg.py:18-87 (`Ckg`, `var_Ckg`)."""
from functools import lru_cache

import numpy as np

from lbgcmb_b200.Cgg import Cgg, Ngg, _each
from lbgcmb_b200.Nkk import Nkk
from lbgcmb_b200.lensing import lensing_kernel, Ckk
from lbgcmb_b200.params import get_params
from lbgcmb_b200.pmh import engine_of
from lbgcmb_b200.quadrature import simps_weights
from lbgcmb_b200.sliced_pz import sliced_pz
from lbgcmb_b200.utils import comoving_distance

params = get_params()


@lru_cache(maxsize=64)
def _ckg_nodes(zmin, zmax):
    dz = 0.01
    zs = np.arange(zmin, zmax, dz)
    out = (zs, comoving_distance(zs), simps_weights(zs.size, zs[1] - zs[0]), lensing_kernel(zs))
    for a in out:
        a.setflags(write=False)
    return out


def ckg_nodes(zmin, zmax):
    """z nodes (no end point, Ckg.py:19-20), chi(z) and Simpson weights (even length -> scipy-1.2.1 'avg' rule); grid
    constants of the fixed cosmology, kept per (zmin, zmax)."""
    return _ckg_nodes(float(zmin), float(zmax))[:3]


def ckg_weights(zs, chis, w, zmin, zmax, survey_pz, bz):
    """Per-node Limber weights of the full z integral (Ckg.py:42-57): quadrature x W_kappa p / chi^2 x b / h^2."""
    cached = _ckg_nodes(float(zmin), float(zmax))
    wk = cached[3] if zs is cached[0] else lensing_kernel(zs)
    prefactor = wk * sliced_pz(zs, zmin, zmax, survey_pz) / chis ** 2.
    return w * prefactor * _each(bz, zs) / params['h_100'] ** 2.


def Ckg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, zeff=True):
    dz = 0.01
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = np.asarray(Llls, dtype=np.float64)
    if zeff:
        zs = np.arange(zmin, zmax, dz)
        ps = survey_pz(zs)
        norm = np.sum(ps) * dz
        ps = ps / norm
        zg = np.sum(ps * zs) * dz / np.sum(dz * ps)                     # Ckg.py:29
        chi_g = comoving_distance(zg)
        amp = float(bz(zg)) * lensing_kernel(zg) / chi_g ** 2. / params['h_100'] ** 2.
        return engine.limber(Ld, [zg], [chi_g], [[amp]], nonlinear=nonlinear)[0]
    zs, chis, w = ckg_nodes(zmin, zmax)
    weights = ckg_weights(zs, chis, w, zmin, zmax, survey_pz, bz)
    return engine.limber(Ld, zs, chis, weights[None, :], nonlinear=nonlinear)[0]


def var_Ckg(Pk_interps, lensCl_interps, nolensCl_interps, Llls, zmin, zmax, survey_pz, bz, nbar, fsky,
            ngg=None, nkk=None, thetab=1., DeltaT=1., zeff=True, modesample_lim=False, iterative=False,
            pickleit=True, rho=None):
    ckk = Ckk(Pk_interps, Llls)
    cgg = Cgg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, zeff=zeff)
    ckg = Ckg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, zeff=zeff)
    if rho is not None:
        ckk = (1. - rho(Llls) ** 2.) * ckk                              # delens by the provided rho(L)
    if modesample_lim:
        ngg = np.zeros_like(Llls)
    else:
        if ngg is None:
            ngg = Ngg(Llls, zmin, zmax, survey_pz, nbar)
    if nkk is None:
        nkk = Nkk(lensCl_interps, nolensCl_interps, Llls, ['TT', 'TE', 'EE', 'EB'], thetab, DeltaT, iterative=iterative,
                  pickleit=pickleit)
    num = (ckk + nkk) * (cgg + ngg) + ckg ** 2.
    den = (2. * Llls + 1.) * fsky
    return num / den
