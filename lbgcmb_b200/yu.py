"""Multi-tracer delensing correlation rho_L; restates the script body of reference yu.py:76-113 as a function.

For every L: kappa-g and g-g spectra of all tracers (batched Limber, one launch per distinct z grid), shot noise on the
diagonal, correlation matrix, rho_L^2 = r^T rho^-1 r with r_i = C^{kappa g_i} / sqrt(C^{kk} C^{g_i g_i}).  The per-L
solves run on the device (lbg_fisher_accum, n_par = 1)."""
import numpy as np

from lbgcmb_b200 import fisher as _fisher
from lbgcmb_b200.Cgg import Cgg, Ngg
from lbgcmb_b200.Ckg import Ckg


def rho_L(Pk_interps, Llls, ckk, samples, zmin=0.01, zmax=10.00, device=None):
    """samples: list of (b(z), p(z), nbar [deg^-2]) as yu.py:59; returns rho_L [len(Llls)] (yu.py:108 `np.sqrt(interim)`)."""
    Llls = np.asarray(Llls)
    nt = len(samples)
    kg = np.array([Ckg(Pk_interps, Llls, zmin, zmax, p, b, zeff=False) for (b, p, n) in samples])              # [nt, nL]
    gg = np.array([[Cgg(Pk_interps, Llls, zmin, zmax, p, b, bz2=b2, survey_pz2=p2, zeff=False) for (b, p, n) in samples]
                   for (b2, p2, n2) in samples])                                                                # [nt, nt, nL]
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
