"""Where the host side of one N0 call goes (run on the GPU box): python scripts/n0_host_time.py"""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import Nkk, _ffi
from lbgcmb_b200.prep_camb import cl_interps_from_tables
L = _ffi.lib()
acc = {}
for name in ('lbg_upload_cl_tables', 'lbg_n0_ho02'):
    fn = getattr(L, name)
    def wrap(*a, _fn=fn, _n=name):
        t = time.perf_counter(); r = _fn(*a); acc[_n] = time.perf_counter() - t; return r
    setattr(L, name, wrap)
li, ni = cl_interps_from_tables(*bench.make_inputs())
Ls = np.arange(50, 5001)
for rep in range(4):
    t = time.perf_counter()
    Nkk.n0_terms(li, ni, Ls, bench.TERMS, 1.4, 1.0, True)
    tot = time.perf_counter() - t
print('n0_terms %.2f ms: upload %.2f, lbg_n0_ho02 %.2f (kernels %.2f, main %.2f), python %.2f' % (
    tot * 1e3, acc['lbg_upload_cl_tables'] * 1e3, acc['lbg_n0_ho02'] * 1e3, Nkk.last_stats['kernel_ms'], Nkk.last_stats['main_kernel_ms'],
    (tot - acc['lbg_upload_cl_tables'] - acc['lbg_n0_ho02']) * 1e3))
