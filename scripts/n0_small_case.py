"""Tiny N0 call (3 L) for compute-sanitizer / debugging: python scripts/n0_small_case.py [fold_phi: 0 | 1 | 2]"""
import os
import sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
from lbgcmb_b200.synthetic import cmb_cl_tables

fold = int(sys.argv[1]) if len(sys.argv) > 1 else 1
ell, lens, nolens = cmb_cl_tables()
li, ni = cl_interps_from_tables(ell, lens, nolens)
Ls = np.array([50, 3000, 5000])
out, _ = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=(2 if fold == 2 else bool(fold)))
print(out, 'kernel path', Nkk.last_stats['path'])
