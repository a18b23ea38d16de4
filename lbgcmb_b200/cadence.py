"""LSST-lens clustering-redshift forecast; mirrors reference cadence/cadence.py:43-191 (`lenses_bz`, `Cij`, `Fisher`).

Host: the per-shell scalars (cadence.py:67-95).  Device: one thin-shell Limber C_gg per shell (Cgg.Cgg, zeff=True), then
A00(l), S(l), the Schur-Limber diagonal and the n_z x n_z Fisher sums over the multipoles inside the reference's L cut
(csrc/fisher.cu through lbg_cadence_fisher).  The final inverse is numpy, as in the reference (:168).

Faithful to the function as written, including
  * `dz` being a module global the reference only defines in its __main__ block (:209) -- here a module attribute;
  * LCUT being whatever the LAST shell of the first loop left behind (:70-71, used at :140,157,163);
  * the running `Interim` sum of :156-165 (each Fisher[i, j] receives base_i + sum_{j' <= j} cross_ij').
"""
import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200.Cgg import Cgg
from lbgcmb_b200.growth_rate import growth_factor
from lbgcmb_b200.pmh import Pmm
from lbgcmb_b200.pz2nbar import nbar_convert
from lbgcmb_b200.utils import comoving_distance
from lbgcmb_b200.whitebook_pz import lsst_chang_pz

KMAX = 0.3        # eqn. (4) of D1.1, pg. 46 of the LSST white book (cadence.py:23)
dz = 0.1          # cadence.py:209

last = {}


def lenses_bz(zee):
    aa = 1. / (1. + zee)
    return 0.95 / growth_factor(aa)


def Cij(Pk_interps, Llls, zmin, zmax):
    zee = (zmin + zmax) / 2.
    chi = comoving_distance(zee)
    dchi = comoving_distance(zmax) - comoving_distance(zmin)
    ks = (Llls + 0.5) / chi
    return Pmm(Pk_interps, ks, zee, 'nlinear') / chi ** 2. / dchi


def Fisher(Pk_interps, Llls, zs, tNs, tNp, pz, bz, fsky=0.1, fover=0.0, sources=False, printit=True):
    """Returns rows [z, Ns, Np, Schur-Limber % error, Limber % error] (cadence.py:178-191)."""
    zs = np.asarray(zs, dtype=np.float64)
    Ld = np.asarray(Llls, dtype=np.float64)
    bins = np.zeros((zs.size, 8))
    cggs = np.empty((zs.size, Ld.size))
    LMAX = None
    for i, zee in enumerate(zs):
        LMAX = np.ceil(KMAX * comoving_distance(zee) - 0.5).astype('int')
        bs = bz(zee)
        Ns = nbar_convert(tNs, unit='str') * pz(zee) * dz
        bp = 1. + zee if sources else lenses_bz(zee)
        Np = nbar_convert(tNp, unit='str') * lsst_chang_pz(zee, ilim=25.3, source=False) * dz
        wp, ws = Np, Ns
        wps = fover * np.min([wp, ws])
        bins[i, :7] = [bp, bs, Np, Ns, wp, ws, wps]
        cggs[i] = Cgg(Pk_interps, Llls, zee - dz / 2., zee + dz / 2., pz, bz,
                      survey_pz2=lambda z: lsst_chang_pz(z, ilim=25.3, source=sources), bz2=lenses_bz, zeff=True)
    w = np.where(Ld <= LMAX, 2. * Ld + 1., 0.0)                        # LCUT of the last shell
    ctx = _ffi.context()
    F, Fsl, S = np.empty((zs.size, zs.size)), np.empty(zs.size), np.empty(Ld.size)
    _ffi.check(_ffi.lib().lbg_cadence_fisher(ctx.handle, Ld.size, zs.size, _ffi._dp(_ffi.f64(cggs)), _ffi._dp(_ffi.f64(bins)),
                                             _ffi._dp(_ffi.f64(w)), float(fsky), _ffi._dp(F), _ffi._dp(Fsl), _ffi._dp(S)))
    iFish = np.linalg.inv(F)
    diFish = np.diag(iFish)
    Np, Ns = bins[:, 2], bins[:, 3]
    ferr_ii = np.sqrt(1. / Fsl) / Np
    last.update(Fisher=F, Fii=Fsl, S=S, bins=bins, cgg=cggs, LMAX=int(LMAX), ferr_ii=ferr_ii, iFisher=iFish)
    if printit:
        for i, zz in enumerate(zs):
            dstr = "\tz:  %.2lf \t\t Schur-Limber:  %.2lf \t\t Limber:  %.2lf" % (zz, 100. * ferr_ii[i], 100. * np.sqrt(diFish[i]) / Np[i])
            dstr += "\t\t Schur at Lmin: %.2lf, and Lmax: %.2lf" % (S[0], S[-1])
            print(dstr)
    output = []
    for outz in zs:
        index = np.where(np.abs((zs - outz)) == np.min(np.abs(zs - outz)))[0]
        k = index[0]
        output.append([outz, Ns[k], Np[k], 100. * ferr_ii[k], 100. * np.sqrt(diFish[k]) / Np[k]])
    output = np.array(output)
    if printit:
        print(output)
    return output
