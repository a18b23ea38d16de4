"""Multi-tracer delensing correlation rho_L; restates the script body of reference yu.py:76-113 as a function.

For every L: kappa-g and g-g spectra of all tracers (two batched Limber launches: n_t and n_t^2 entries), shot noise on the
diagonal, correlation matrix, rho_L^2 = r^T rho^-1 r with r_i = C^{kappa g_i} / sqrt(C^{kk} C^{g_i g_i}).  The per-L
solves run on the device (lbg_fisher_accum, n_par = 1)."""
import numpy as np

from lbgcmb_b200 import fisher as _fisher
from lbgcmb_b200.Cgg import Ngg, cgg_nodes, cgg_weights
from lbgcmb_b200.Ckg import ckg_nodes, ckg_weights
from lbgcmb_b200.pmh import engine_of


def rho_L(Pk_interps, Llls, ckk, samples, zmin=0.01, zmax=10.00, device=None):
    """samples: list of (b(z), p(z), nbar [deg^-2]) as yu.py:59; returns rho_L [len(Llls)] (yu.py:108 `np.sqrt(interim)`)."""
    Llls = np.asarray(Llls)
    nt = len(samples)
    # what Ckg.Ckg / Cgg.Cgg (zeff=False) return per tracer and per tracer pair, from two batched Limber launches
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = Llls.astype(np.float64)
    zs, chis, w = ckg_nodes(zmin, zmax)
    wk = np.stack([ckg_weights(zs, chis, w, zmin, zmax, p, b) for (b, p, n) in samples])
    kg = engine.limber(Ld, zs, chis, wk, nonlinear=nonlinear)                                                    # [nt, nL]
    zs, chis, w = cgg_nodes(zmin, zmax)
    wg = np.stack([cgg_weights(zs, chis, w, zmin, zmax, p, b, p2, b2) for (b2, p2, n2) in samples for (b, p, n) in samples])
    gg = engine.limber(Ld, zs, chis, wg, nonlinear=nonlinear).reshape(nt, nt, Llls.size)                         # [nt, nt, nL]
    for i, (b, p, n) in enumerate(samples):
        gg[i, i] += Ngg(Llls.astype(np.float64), zmin, zmax, p, n)                                              # yu.py:89-91
    diag = np.array([gg[i, i] for i in range(nt)])                                                              # [nt, nL]
    rho = gg / (np.sqrt(diag)[:, None, :] * np.sqrt(diag)[None, :, :])                                          # :96-98
    r = kg / (np.sqrt(np.asarray(ckk, dtype=np.float64))[None, :] * np.sqrt(diag))                              # :100
    cov = np.ascontiguousarray(np.transpose(rho, (2, 0, 1)))
    dD = np.ascontiguousarray(np.transpose(r, (1, 0))[:, None, :])
    _, _, Q = _fisher.fisher_accum(cov, dD, np.ones(Llls.size), per_ell=True, device=device)
    return np.sqrt(Q[:, 0, 0])


def save_rho(path, Llls, rho):
    """The rho/*.txt format of yu.py:113: columns L, 1 - rho_L^2."""
    np.savetxt(path, np.c_[Llls, (1. - rho ** 2.)])
