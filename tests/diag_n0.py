"""GPU diagnostic: per-L, per-estimator relative error of lbg_n0_ho02 vs the numpy oracle and golden vectors."""
import glob, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from oracle import n0 as on0

for f in sorted(glob.glob(os.path.join(ROOT, 'tests/golden/n0_root_*.npz'))):
    g = np.load(f)
    li, ni = cl_interps_from_tables(g['ell'], g['lensed'], g['unlensed'])
    for fold in (False, True):
        out, _ = Nkk.n0_terms(li, ni, g['Ls'], [str(t) for t in g['terms']], float(g['thetab']), float(g['DeltaT']),
                              bool(g['iterative']), fold_phi=fold, Tcmb=float(g['Tcmb']), cos_sin=(g['cos_phi'], g['sin_phi']))
        print(os.path.basename(f), 'fold', fold, out / g['nkk_reference'] - 1)

g = np.load(os.path.join(ROOT, 'tests/golden/n0_root_cmbs4.npz'))
ell, lens, nolens = g['ell'], g['lensed'], g['unlensed']
li, ni = cl_interps_from_tables(ell, lens, nolens)
oi = (on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens))
Ls = np.array([int(x) for x in sys.argv[1:]] or [50, 3000])
for terms in (['TT'], ['TE'], ['EE'], ['EB']):
    ref = on0.nkk_root(oi[0], oi[1], Ls, tuple(terms), 1.4, 1.0, False)
    for fold in (False, True):
        got = Nkk.n0_terms(li, ni, Ls, terms, 1.4, 1.0, False, fold_phi=fold)[0]
        print(terms, 'fold', fold, got / ref - 1)
