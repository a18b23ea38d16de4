"""CMB lensing kernel, C_kk and its variance; mirrors reference cmb/lensing.py:14-66."""
from functools import lru_cache

import numpy as np

from lbgcmb_b200.cosmo import cosmo, C_KMS
from lbgcmb_b200.params import get_params
from lbgcmb_b200.pmh import engine_of
from lbgcmb_b200.quadrature import simps_weights
from lbgcmb_b200.utils import comoving_distance

params = get_params()


def lensing_kernel(z):
    chi = comoving_distance(z)
    chi_scatter = comoving_distance(params['zscatter'])
    return (3. / 2.) * params['om_m'] * (100. * params['h_100'] / C_KMS) ** 2. * (1. + z) * chi * (1. - chi / chi_scatter)


@lru_cache(maxsize=8)
def _ckk_nodes(zmax):
    """z nodes, chi(z) and kernel x Simpson weights of cmb/lensing.py:27-43: grid constants of the fixed cosmology."""
    z = np.linspace(0.001, zmax, 2000)
    chis = comoving_distance(z)
    prefactor = (C_KMS / cosmo.H(z).value) * (lensing_kernel(z) / chis) ** 2.
    weights = simps_weights(z.size, z[1] - z[0]) * prefactor / params['h_100'] ** 3.
    for a in (z, chis, weights):
        a.setflags(write=False)
    return z, chis, weights


def Ckk(Pk_interps, Llls, pickle=False, zmax=params['zscatter']):
    """C_kk(L) in the extended Limber approximation; cmb/lensing.py:21-52 (2000 z nodes to z*, Simpson)."""
    z, chis, weights = _ckk_nodes(float(zmax))
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    result = engine.limber(np.asarray(Llls, dtype=np.float64), z, chis, weights[None, :], nonlinear=nonlinear)[0]
    if pickle:
        from pickle import dump
        dump(result, open("pickle/kk.p", "wb"))
    return result


def var_Ckk(Llls, fsky, nkk, Pk_interps=None, samplevar_lim=False):
    ckk = Ckk(Pk_interps, Llls)
    if samplevar_lim:
        num = ckk
    else:
        num = ckk + nkk
    num = 2. * num ** 2.
    den = (2. * Llls + 1.) * fsky
    return num / den
