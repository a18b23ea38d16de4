"""cProfile of the host side of bench.py's forecast stage (run on the GPU box): python scripts/profile_forecast.py"""
import cProfile
import os
import pstats
import sys
import time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import interloper, pmh, snr, _ffi

kk, pp, bands = bench.forecast_inputs()
Ls = np.arange(50, 5001).astype(np.float64)
nkk = 1e-8 * np.ones_like(Ls)


T = pmh.PkTable(kk, pp)
Pk = pmh.get_PkInterps(T)


def run():
    T.engine.invalidate_device()
    return bench.forecast_stage((interloper, snr), Pk, bands, Ls, np.ones_like(Ls), nkk)


for _ in range(2):
    t = time.perf_counter(); run(); print('stage s', time.perf_counter() - t)
ctx = _ffi.context()
ctx.kernel_totals(reset=True)
pr = cProfile.Profile(); pr.enable(); run(); pr.disable()
print('kernel totals', ctx.kernel_totals())
pstats.Stats(pr).sort_stats('tottime').print_stats(28)

# host side of N0
from lbgcmb_b200 import Nkk
from lbgcmb_b200.prep_camb import cl_interps_from_tables
ell, lens, nolens = bench.make_inputs()
li, ni = cl_interps_from_tables(ell, lens, nolens)
Li = np.arange(50, 5001)
for _ in range(2):
    Nkk.n0_terms(li, ni, Li, bench.TERMS, 1.4, 1.0, True)
pr = cProfile.Profile(); pr.enable(); Nkk.n0_terms(li, ni, Li, bench.TERMS, 1.4, 1.0, True); pr.disable()
print('n0 kernel ms', Nkk.last_stats)
pstats.Stats(pr).sort_stats('cumulative').print_stats(20)
