"""Multi-level N0 case for ncu (every 10th L of the bench grid, six noise levels of the 1.4' beam -> two launches of the 3-level kernel):
python scripts/n0_levels_profile_case.py"""
import json
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from lbgcmb_b200.synthetic import cmb_cl_tables

li, ni = cl_interps_from_tables(*cmb_cl_tables())
Ls = np.arange(50, 5001, 10)
dts = [1., 2., 5., 10., 20., 30.]
Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, False)
single = dict(Nkk.last_stats)
for rep in range(2):
    out, _ = Nkk.n0_levels(li, ni, Ls, 1.4, dts)
s = Nkk.last_stats
print(json.dumps({'case': 'n0_levels_profile_case', 'n_L': int(Ls.size), 'n_lev': len(dts), 'main_kernel_ms': s['main_kernel_ms'],
                  'ms_per_level': s['main_kernel_ms'] / len(dts), 'single_level_kernel_ms': single['main_kernel_ms'],
                  'nodes_unmasked': single['nodes_unmasked'], 'path': s['path'], 'n0_first': float(out[0, 0])}))
