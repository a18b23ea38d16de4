"""One lbg_field_fisher call at the configs[3] shape (121 fields, arrowhead pattern, 8 parameters) on 1480 multipoles, for ncu:
  ncu --set full --clock-control none --import-source on -k regex:field_fisher -c 1 -o gpurun_out/ff python scripts/field_fisher_profile_case.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import _ffi                                    # noqa: E402
from lbgcmb_b200.fisher import field_fisher                     # noqa: E402

n_f, n_ell, n_par = 121, 1480, 8
rng = np.random.default_rng(0)
spectra = [(0, 0)] + [(0, i) for i in range(1, n_f)] + [(i, i) for i in range(1, n_f)]
d = 10. ** rng.uniform(-8, -4, (n_f, n_ell))
tot = np.concatenate([d[:1], 0.05 * rng.uniform(-1, 1, (n_f - 1, n_ell)) * np.sqrt(d[:1] * d[1:]), d[1:]])
dC = tot[None] * rng.uniform(-2, 2, (n_par, len(spectra), 1))
w = np.ones(n_ell)
ctx = _ffi.context()
for _ in range(3):
    t0 = time.perf_counter()
    F = field_fisher(n_f, spectra, tot, dC, w)
    t1 = time.perf_counter()
    print('kernel ms %.3f  wall ms %.3f' % (ctx.last_kernel_ms()[0], (t1 - t0) * 1e3))
assert np.all(np.linalg.eigvalsh(F) > 0)
