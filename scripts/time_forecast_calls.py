"""Wall time of every library call inside bench.py's forecast stage (run on the GPU box)."""
import os, sys, time, collections
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import interloper, pmh, snr, _ffi

L = _ffi.lib()
acc = collections.OrderedDict()
for name in ('lbg_limber', 'lbg_pk_build_table', 'lbg_snr', 'lbg_gaussian_cov', 'lbg_fisher_accum', 'lbg_upload_pk_linear', 'lbg_halofit_sigma2'):
    fn = getattr(L, name)
    def wrap(*a, _fn=fn, _n=name):
        t = time.perf_counter(); r = _fn(*a); acc.setdefault(_n, []).append(time.perf_counter() - t); return r
    setattr(L, name, wrap)
kk, pp, bands = bench.forecast_inputs()
Ls = np.arange(50, 5001).astype(np.float64)
nkk = 1e-8 * np.ones_like(Ls)
T = pmh.PkTable(kk, pp); Pk = pmh.get_PkInterps(T)
for rep in range(3):
    acc.clear()
    T.engine.invalidate_device()
    t = time.perf_counter()
    bench.forecast_stage((interloper, snr), Pk, bands, Ls, np.ones_like(Ls), nkk)
    tot = time.perf_counter() - t
print('stage %.2f ms' % (tot * 1e3))
for k, v in acc.items():
    print('%-22s n=%3d total %.2f ms  max %.2f ms' % (k, len(v), sum(v) * 1e3, max(v) * 1e3))
print('library total %.2f ms' % (sum(sum(v) for v in acc.values()) * 1e3))
