"""Clustering-redshift N_p(z_i) Fisher forecast (McQuinn & White); mirrors reference mqw/py/mqw.py:18-127 (`Fisher`).

Host: the per-shell scalars (mqw.py:24-41).  Device: the shared Limber C_gg (one call; the reference repeats the identical
call for each of the 20 shells, mqw.py:48), A00(l), S(l) and the n_z x n_z Fisher sums over l (csrc/fisher.cu
mqw_schur_kernel / mqw_fisher_kernel through lbg_mqw_fisher).  The final inverse is numpy, as in the reference (:106)."""
import numpy as np

from lbgcmb_b200 import fisher as _fisher
from lbgcmb_b200.Cgg import Cgg
from lbgcmb_b200.pz2nbar import nbar_convert

last = {}     # Fisher matrix, Schur-Limber diagonal, S(l) and per-shell numbers of the latest call


def Fisher(Pk_interps, Llls, tNs, tNp, ps, _bs, pp, _bp, dz=0.1, fsky=0.1, zmin=3.0, zmax=4.0, fover=0.0, percentiles=[],
           intlp_zs=[], printit=False, cgg=None):
    """Returns [Schur-Limber %, Limber %] fractional errors on N_p(z) at each requested redshift percentile."""
    zs = np.arange(zmin + dz / 2., zmax + dz / 2., dz)
    bins = np.zeros((zs.size, 7))
    for i, zee in enumerate(zs):
        bs = _bs(zee)
        bp = _bp(zee)
        Ns = nbar_convert(tNs, unit='str') * ps(zee) * dz
        Np = nbar_convert(tNp, unit='str') * pp(zee) * dz
        wp, ws = Np, Ns                                                # eqn. (11) of MQW16, fsat = 0
        wps = fover * np.min([wp, ws])
        bins[i] = [bp, bs, Np, Ns, wp, ws, wps]
    if cgg is None:
        ub = lambda z: 1.0
        cgg = Cgg(Pk_interps, Llls, zmin, zmax, pp, ub, survey_pz2=ps, bz2=ub, zeff=True)     # mqw.py:47-48
    Ld = np.asarray(Llls, dtype=np.float64)
    F, Fsl, S = _fisher.mqw_fisher(cgg, bins, 2. * Ld + 1., fsky)
    iFish = np.linalg.inv(F)
    diFish = np.diag(iFish)
    Np = bins[:, 2]
    ferr_ii = np.sqrt(1. / Fsl) / Np                                    # mqw.py:84-85
    last.update(Fisher=F, Fii=Fsl, S=S, zs=zs, bins=bins, cgg=np.asarray(cgg), ferr_ii=ferr_ii, iFisher=iFish)
    if printit:
        for i, zz in enumerate(zs):
            dstr = "\tz:  %.2lf \t\t Schur-Limber:  %.2lf \t\t Limber:  %.2lf" % (zz, 100. * ferr_ii[i], 100. * np.sqrt(diFish[i]) / Np[i])
            dstr += "\t\t Schur at Lmin: %.2lf, and Lmax: %.2lf" % (S[0], S[-1])
            print(dstr)
    output = []
    for outz in percentiles:
        index = np.where(np.abs((zs - outz)) == np.min(np.abs(zs - outz)))[0]
        output += [100. * ferr_ii[index[0]], 100. * np.sqrt(diFish[index]) / Np[index[0]]]
    return output
