"""GPU diagnostic (a checker): EE integral of lbg_n0_ho02 on a sub-grid of l1 rows around the end of a short C_l table, against a
direct numpy evaluation of cmb/Nkk.py:66-133 with scipy interp1d."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from scipy.interpolate import interp1d
from lbgcmb_b200 import Nkk
from lbgcmb_b200.DetectorNoise import DetectorNoise
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from lbgcmb_b200.synthetic import cmb_cl_tables
ell, lens, nolens = cmb_cl_tables()
n = 4001
L = 60.0
li, ni = cl_interps_from_tables(ell[:n], lens[:, :n], nolens[:, :n])
fE = interp1d(ell[:n], nolens[1][:n], kind='cubic', bounds_error=False, fill_value=0.0)
fEl = interp1d(ell[:n], lens[1][:n], kind='cubic', bounds_error=False, fill_value=0.0)
g_ell, g_wl, pphi, wphi = Nkk.root_grids()
for lo, hi in ((3000.5, 3900.), (3900.5, 3990.), (3990.5, 4000.), (4000.5, 4010.), (4010.5, 4100.), (4100.5, 4999.5)):
    sel = (g_ell >= lo) & (g_ell <= hi)
    e, w = g_ell[sel], g_wl[sel]
    got = Nkk.n0_terms(li, ni, np.array([L]), ('EE',), 1.4, 7.0, False, grids=(e, w, pphi, wphi), exact_u_below=0.0, fold_phi=False)[1]['EE'][0]
    ll, pp = np.meshgrid(e, pphi, indexing='ij')
    ux, uy = ll * np.cos(pp), ll * np.sin(pp)
    vx, vy = L - ux, -uy
    nu, nv = (ux ** 2 + uy ** 2) ** 0.5, (vx ** 2 + vy ** 2) ** 0.5
    c2 = np.cos(2. * (np.arctan2(uy, ux) - np.arctan2(vy, vx)))
    f = (fE(nu) * L * ux + fE(nv) * L * vx) * c2
    Tu = fEl(nu) + DetectorNoise(nu, 1.4, 7.0, 'EE')
    Tv = fEl(nv) + DetectorNoise(nv, 1.4, 7.0, 'EE')
    m = (nu <= 5000) & (nv <= 5000) & (nv >= 2)
    integ = f * f / (2. * Tu * Tv) * m
    want = np.sum(w[:, None] * e[:, None] * wphi[None, :] * integ)
    print('rows %.1f..%.1f  lib %.6e  numpy %.6e  rel %.2e' % (lo, hi, got, want, got / want - 1 if want else 0))
