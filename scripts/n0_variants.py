"""Time the N0 kernel variants on the bench workload and check them against each other (run on the GPU box)."""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from lbgcmb_b200.synthetic import cmb_cl_tables

ell, lens, nolens = cmb_cl_tables()
li, ni = cl_interps_from_tables(ell, lens, nolens)
Ls = np.arange(50, 5001)
res = {}
for v in sys.argv[1:] or ['0', '1']:
    os.environ['LBG_N0_VARIANT'] = v
    for rep in range(3):
        out, _ = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True)
    res[v] = out
    print('variant', v, 'kernel_ms %.2f' % Nkk.last_stats['kernel_ms'], 'launches', Nkk.last_stats['launches'], flush=True)
base = res[sorted(res)[0]]
for v in res:
    print('variant', v, 'max rel diff vs first: %.3e' % np.max(np.abs(res[v] / base - 1)))
