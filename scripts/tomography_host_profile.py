"""cProfile of the second (warm) tomographic_fisher call of scripts/tomography_fisher_bench.py's workload: where the host time goes."""
import cProfile
import os
import pstats
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import scripts.tomography_fisher_bench as tb          # noqa: E402
from lbgcmb_b200 import tomography as tm              # noqa: E402

Lf = np.arange(50., 5001.)
tracers = []
for name, (z0, sig, nbar, b0) in tb.TRACERS.items():
    lo0 = max(z0 - tb.N_BIN * tb.DZ / 2., 0.05)
    tracers.append({'name': name, 'edges': lo0 + tb.DZ * np.arange(tb.N_BIN + 1),
                    'dndz': (lambda z0_, s_: (lambda z: np.exp(-0.5 * ((np.asarray(z) - z0_) / s_) ** 2.)))(z0, sig),
                    'bz': (lambda b_, z_: (lambda z: b_ * (1. + z) / (1. + z_)))(b0, z0), 'nbar': 1e6})
nkk = 2e-8 * np.ones_like(Lf)
tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), 0.4, nkk)
pr = cProfile.Profile()
pr.enable()
tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), 0.4, nkk)
pr.disable()
pstats.Stats(pr).sort_stats('cumtime').print_stats(45)
