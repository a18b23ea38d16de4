"""Small N0 case for ncu (every 10th L of the bench grid, same grids / beam / noise): python scripts/n0_profile_case.py
Prints the node counts scripts/ncu_n0_json.py needs (the kernel counts its own unmasked nodes)."""
import json
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from lbgcmb_b200.synthetic import cmb_cl_tables

ell, lens, nolens = cmb_cl_tables()
li, ni = cl_interps_from_tables(ell, lens, nolens)
Ls = np.arange(50, 5001, 10)
for rep in range(2):
    out, _ = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True)
s = Nkk.last_stats
print(json.dumps({'case': 'n0_profile_case', 'n_L': int(Ls.size), 'kernel_ms': s['kernel_ms'], 'main_kernel_ms': s['main_kernel_ms'],
                  'nodes_walked': s['nodes'], 'nodes_unmasked': s['nodes_unmasked'], 'path': s['path'], 'n0_first': float(out[0])}))
