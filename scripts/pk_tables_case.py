"""The P(k,z) table build of one forecast step (three z grids, one library call), four times: for ncu launch lists."""
import os
import sys
import time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import pmh
from lbgcmb_b200.Cgg import _cgg_nodes
from lbgcmb_b200.Ckg import _ckg_nodes
from lbgcmb_b200.lensing import _ckk_nodes
kk, pp, bands = bench.forecast_inputs()
T = pmh.PkTable(kk, pp)
grids = [_cgg_nodes(0.01, 10.0)[0], _ckg_nodes(0.01, 10.0)[0], _ckk_nodes(1100.)[0]]
for rep in range(4):
    T.engine.invalidate_device()
    t0 = time.perf_counter()
    T.engine.build_tables(grids)
    print('build_tables %.3f ms' % ((time.perf_counter() - t0) * 1e3), 'kernel ms', T.engine._ctx().last_kernel_ms())
