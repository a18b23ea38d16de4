"""Limber C_kg and its variance; mirrors reference Ckg.py:18-87 (`Ckg`, `var_Ckg`)."""
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


def Ckg(Pk_interps, Llls, zmin, zmax, survey_pz, bz, zeff=True):
    dz = 0.01
    zs = np.arange(zmin, zmax, dz)
    ps = survey_pz(zs)
    norm = np.sum(ps) * dz
    ps = ps / norm
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = np.asarray(Llls, dtype=np.float64)
    if zeff:
        zg = np.sum(ps * zs) * dz / np.sum(dz * ps)                     # Ckg.py:29
        chi_g = comoving_distance(zg)
        amp = float(bz(zg)) * lensing_kernel(zg) / chi_g ** 2. / params['h_100'] ** 2.
        return engine.limber(Ld, [zg], [chi_g], [[amp]], nonlinear=nonlinear)[0]
    chis = comoving_distance(zs)
    prefactor = lensing_kernel(zs) * sliced_pz(zs, zmin, zmax, survey_pz) / chis ** 2.
    weights = simps_weights(zs.size, zs[1] - zs[0]) * prefactor * _each(bz, zs) / params['h_100'] ** 2.
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
