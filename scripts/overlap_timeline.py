"""Timeline of one overlapped step of bench.py (run on the GPU box): when each library call starts / ends on each thread."""
import os, sys, time, threading
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import Nkk, interloper, pmh, snr, _ffi
from lbgcmb_b200.prep_camb import cl_interps_from_tables

L = _ffi.lib()
log = []
T0 = [0.0]
for name in ('lbg_limber', 'lbg_pk_build_table', 'lbg_snr', 'lbg_gaussian_cov', 'lbg_fisher_accum', 'lbg_upload_pk_linear',
             'lbg_halofit_sigma2', 'lbg_n0_ho02', 'lbg_upload_cl_tables'):
    fn = getattr(L, name)
    def wrap(*a, _fn=fn, _n=name):
        t = time.perf_counter(); r = _fn(*a)
        log.append((threading.current_thread().name[:8], _n, (t - T0[0]) * 1e3, (time.perf_counter() - T0[0]) * 1e3)); return r
    setattr(L, name, wrap)
ell, lens, nolens = bench.make_inputs()
li, ni = cl_interps_from_tables(ell, lens, nolens)
kk, pp, bands = bench.forecast_inputs()
Ls = np.arange(50, 5001); Lf = Ls.astype(np.float64)
T = pmh.PkTable(kk, pp); Pk = pmh.get_PkInterps(T)
for rep in range(4):
    log.clear()
    T0[0] = time.perf_counter()
    fut = Nkk.n0_terms_async(li, ni, Ls, bench.TERMS, 1.4, 1.0, True)
    T.engine.invalidate_device()
    rows = bench.forecast_stage((interloper, snr), Pk, bands, Lf, np.ones_like(Lf), fut)
    tot = (time.perf_counter() - T0[0]) * 1e3
print('step %.2f ms' % tot)
for th, n, a, b in sorted(log, key=lambda r: r[2]):
    print('%-9s %-22s %7.2f -> %7.2f  (%.2f)' % (th, n, a, b, b - a))
