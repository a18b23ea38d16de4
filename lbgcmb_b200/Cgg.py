"""Limber C_gg, shot noise and variance; mirrors reference Cgg.py:17-95 (`Cgg`, `var_Cgg`, `Ngg`).

Host (numpy, <= 1000 doubles, as in the reference): the z grid, p(z) normalisation, chi(z), H(z), `sliced_pz`, b(z).
Device (csrc/limber.cu): P_nl(k, z) on every z node (growth + HALOFIT) and the sum over z for every l."""
from functools import lru_cache

import numpy as np

from lbgcmb_b200.cosmo import cosmo, C_KMS
from lbgcmb_b200.params import get_params
from lbgcmb_b200.pmh import engine_of
from lbgcmb_b200.pz2nbar import pz_slice, nbar_convert
from lbgcmb_b200.quadrature import simps_weights
from lbgcmb_b200.sliced_pz import sliced_pz
from lbgcmb_b200.utils import comoving_distance

params = get_params()


def _each(f, zs):
    """b(z) on every node of the z grid.  The reference calls it one node at a time inside its z loop (Cgg.py:58); an
    array-capable callable gives the same numbers in one call, anything else (scalar-only branches such as
    interloper.py:156, np.vectorize'd tables that return the wrong shape) is sampled node by node as the reference does."""
    try:
        v = np.asarray(f(zs), dtype=np.float64)
        if v.shape == zs.shape:
            return v
    except Exception:
        pass
    return np.array([float(f(z)) for z in zs])


@lru_cache(maxsize=64)
def _cgg_nodes(zmin, zmax):
    dz = 0.01
    zs = np.arange(zmin, zmax + dz, dz)
    out = (zs, comoving_distance(zs), simps_weights(zs.size, dz), cosmo.H(zs).value / C_KMS)
    for a in out:
        a.setflags(write=False)
    return out


def cgg_nodes(zmin, zmax):
    """z nodes, chi(z) and scipy-1.2.1 Simpson weights of Cgg.py:18-19,68 (grid constants of the fixed cosmology: kept per
    (zmin, zmax))."""
    return _cgg_nodes(float(zmin), float(zmax))[:3]


def cgg_weights(zs, chis, w, zmin, zmax, survey_pz, bz, survey_pz2=None, bz2=None):
    """Per-node Limber weights of the full z integral (Cgg.py:53-68): quadrature weight x (H/c) p1 p2 / chi^2 x b1 b2 / h."""
    if bz2 is None:
        bz2 = bz
    if survey_pz2 is None:
        survey_pz2 = survey_pz
    p1 = sliced_pz(zs, zmin, zmax, survey_pz)
    p2 = p1 if survey_pz2 is survey_pz else sliced_pz(zs, zmin, zmax, survey_pz2)
    b1 = _each(bz, zs)
    b2 = b1 if bz2 is bz else _each(bz2, zs)
    prefactor = (cosmo.H(zs).value / C_KMS) * p1 * p2 / chis ** 2.
    return w * prefactor * b1 * b2 / params['h_100']


def Cgg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, survey_pz2=None, bz2=None, zeff=True):
    dz = 0.01
    zs, chis, w = cgg_nodes(zmin, zmax)
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = np.asarray(Llls, dtype=np.float64)
    if zeff:
        if bz2 is None:
            bz2 = bz
        if survey_pz2 is None:
            survey_pz2 = survey_pz
        ps = survey_pz(zs)
        norm = np.sum(ps) * dz
        ps = ps / norm
        prefactor = (cosmo.H(zs).value / C_KMS) * sliced_pz(zs, zmin, zmax, survey_pz) \
            * sliced_pz(zs, zmin, zmax, survey_pz2) / chis ** 2.
        zz = dz * np.sum(ps * zs) / (np.sum(ps) * dz)                   # Cgg.py:46
        amp = float(bz(zz)) * float(bz2(zz)) * np.sum(w * prefactor) / params['h_100']
        return engine.limber(Ld, [zz], [comoving_distance(zz)], [[amp]], nonlinear=nonlinear)[0]
    weights = cgg_weights(zs, chis, w, zmin, zmax, survey_pz, bz, survey_pz2, bz2)
    return engine.limber(Ld, zs, chis, weights[None, :], nonlinear=nonlinear)[0]


def Ngg(Llls, zmin, zmax, survey_pz, nbar):
    """Poisson noise 1 / (angular number density); Cgg.py:85-95."""
    spz = pz_slice(zmin, zmax, survey_pz)
    nbar = nbar_convert(nbar, unit='str')
    return np.ones_like(Llls) / (nbar * spz)


def var_Cgg(Llls, zmin, zmax, survey_pz, bz, nbar, fsky, cgg=None, Pk_interps=None, samplevar_lim=False):
    if cgg is None:
        cgg = Cgg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, zeff=True)
    if samplevar_lim:
        num = cgg
    else:
        num = cgg + Ngg(Llls, zmin, zmax, survey_pz, nbar)
    num = 2. * num ** 2.
    den = (2. * Llls + 1.) * fsky
    return num / den
