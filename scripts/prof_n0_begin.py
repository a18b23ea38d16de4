import cProfile, pstats, sys, os, time
sys.path.insert(0, '/root/repo')
import numpy as np
import bench
from lbgcmb_b200 import Nkk, _ffi
from lbgcmb_b200.prep_camb import cl_interps_from_tables
li, ni = cl_interps_from_tables(*bench.make_inputs())
Ls = np.arange(50, 5001)[::8]
for _ in range(5):
    Nkk.n0_begin(li, ni, Ls, bench.TERMS, 1.4, 1.0, True, lane=1).result()
ts = []
for _ in range(30):
    t0 = time.perf_counter(); c = Nkk.n0_begin(li, ni, Ls, bench.TERMS, 1.4, 1.0, True, lane=1); t1 = time.perf_counter(); c.result(); t2 = time.perf_counter()
    ts.append((t1 - t0, t2 - t1))
ts = np.array(ts) * 1e6
print('n0_begin us median', np.median(ts[:, 0]), 'result us', np.median(ts[:, 1]))
pr = cProfile.Profile()
pr.enable()
for _ in range(200):
    Nkk.n0_begin(li, ni, Ls, bench.TERMS, 1.4, 1.0, True, lane=1).result()
pr.disable()
pstats.Stats(pr).sort_stats('tottime').print_stats(14)
