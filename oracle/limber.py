"""CPU restatement of the Limber projections (TEST INFRASTRUCTURE).

root flavour: Cgg.py:17-68 (Cgg), :85-95 (Ngg), Ckg.py:18-59 (Ckg), cmb/lensing.py:14-52 (lensing_kernel, Ckk),
              sliced_pz.py:4-27, pz2nbar.py:6-31, utils.py:33-34, cosmo.py:6.
modi flavour: modi/lensing.py:43-179 (LensingX) -- see oracle/modi_lensing.py.
Pinned against the unmodified reference by oracle/make_golden.py -> tests/golden/limber_root.npz.
"""
import numpy as np

from .simps121 import simps
from .shim.astropy.cosmology import FlatLambdaCDM

C_KMS = 299792.458


class Background(object):
    """params.py:6 + cosmo.py:6 + utils.py:33."""

    def __init__(self, params):
        self.p = params
        self.cosmo = FlatLambdaCDM(H0=100. * params['h_100'], Om0=params['om_m'], Ob0=params['om_b'])

    def chi(self, z):
        return self.p['h_100'] * self.cosmo.comoving_distance(z).value

    def H(self, z):
        return self.cosmo.H(z).value

    def lensing_kernel(self, z):
        chi = self.chi(z)
        chis = self.chi(self.p['zscatter'])
        return (3. / 2.) * self.p['om_m'] * (100. * self.p['h_100'] / C_KMS) ** 2. * (1. + z) * chi * (1. - chi / chis)


def sliced_pz(zz, zmin, zmax, survey_pz):
    dz = 0.001
    zs = np.arange(zmin, zmax, dz)
    ps = survey_pz(zs)
    norm = np.trapezoid(ps, zs)
    result = survey_pz(zz) / norm
    result[zz < zmin] = 0.0
    result[zz > zmax] = 0.0
    dz = list(set(np.around(np.ediff1d(zz), 6)))[0]
    result /= (np.sum(result) * dz)
    return result


def Cgg(bg, Pk, Llls, zmin, zmax, survey_pz, bz, survey_pz2=None, bz2=None, zeff=True):
    dz = 0.01
    zs = np.arange(zmin, zmax + dz, dz)
    ps = survey_pz(zs)
    ps = ps / (np.sum(ps) * dz)
    chis = bg.chi(zs)
    bz2 = bz if bz2 is None else bz2
    survey_pz2 = survey_pz if survey_pz2 is None else survey_pz2
    if zeff:
        zz = dz * np.sum(ps * zs) / (np.sum(ps) * dz)
        ks = (Llls + 0.5) / bg.chi(zz)
        result = Pk['camb_nlin'](zz, ks) * bz(zz) * bz2(zz)
        result = np.broadcast_to(result, (len(zs), len(Llls)))
    else:
        result = np.zeros((len(zs), len(Llls)))
        for i, z in enumerate(zs):
            ks = (Llls + 0.5) / chis[i]
            result[i, :] = Pk['camb_nlin'](z, ks) * bz(z) * bz2(z)
    prefactor = (bg.H(zs) / C_KMS) * sliced_pz(zs, zmin, zmax, survey_pz) * sliced_pz(zs, zmin, zmax, survey_pz2) / chis ** 2.
    integrand = prefactor[:, None] * result
    integrand = integrand / bg.p['h_100']
    return simps(integrand, dx=dz, axis=0)


def Ckg(bg, Pk, Llls, zmin, zmax, survey_pz, bz, zeff=True):
    dz = 0.01
    zs = np.arange(zmin, zmax, dz)
    ps = survey_pz(zs)
    ps = ps / (np.sum(ps) * dz)
    if zeff:
        zg = np.sum(ps * zs) * dz / np.sum(dz * ps)
        chi_g = bg.chi(zg)
        k = (Llls + 0.5) / chi_g
        result = Pk['camb_nlin'](zg, k) * bz(zg) * bg.lensing_kernel(zg) / chi_g ** 2.
        return result / bg.p['h_100'] ** 2.
    chis = bg.chi(zs)
    result = np.zeros((len(zs), len(Llls)))
    for i, z in enumerate(zs):
        ks = (Llls + 0.5) / chis[i]
        result[i, :] = Pk['camb_nlin'](z, ks) * bz(z)
    prefactor = bg.lensing_kernel(zs) * sliced_pz(zs, zmin, zmax, survey_pz) / chis ** 2.
    integrand = prefactor[:, None] * result
    integrand = integrand / bg.p['h_100'] ** 2.
    return simps(integrand, dx=zs[1] - zs[0], axis=0)


def Ckk(bg, Pk, Llls, zmax=None):
    zmax = bg.p['zscatter'] if zmax is None else zmax
    z = np.linspace(0.001, zmax, 2000)
    chis = bg.chi(z)
    result = np.zeros((len(z), len(Llls)))
    for i, zi in enumerate(z):
        ks = (Llls + 0.5) / chis[i]
        result[i, :] = Pk['camb_nlin'](zi, ks)
    prefactor = (C_KMS / bg.H(z)) * (bg.lensing_kernel(z) / chis) ** 2.
    integrand = prefactor[:, None] * result
    integrand = integrand / bg.p['h_100'] ** 3.
    return simps(integrand, dx=z[1] - z[0], axis=0)


def Ngg(Llls, zmin, zmax, survey_pz, nbar):
    zz = np.arange(zmin, zmax, 0.001)
    spz = np.trapezoid(survey_pz(zz), zz, dx=0.001)
    return np.ones_like(Llls) / (nbar / (np.pi / 180.) ** 2. * spz)


def var_auto(Llls, cl, nl, fsky):
    """2 (C + N)^2 / ((2L + 1) fsky): Cgg.py:70-83 (var_Cgg) and cmb/lensing.py:54-66 (var_Ckk); nl = 0 for samplevar_lim."""
    num = cl + nl
    num = 2. * num ** 2.
    den = (2. * Llls + 1.) * fsky
    return num / den


def var_cross(Llls, ckk, nkk, cgg, ngg, ckg, fsky, rho=None):
    """((C_kk (1 - rho^2) + N_kk)(C_gg + N_gg) + C_kg^2) / ((2L + 1) fsky): Ckg.py:60-87 (var_Ckg)."""
    if rho is not None:
        ckk = (1. - rho ** 2.) * ckk
    num = (ckk + nkk) * (cgg + ngg) + ckg ** 2.
    den = (2. * Llls + 1.) * fsky
    return num / den
