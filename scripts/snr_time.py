import sys, time, numpy as np
sys.path.insert(0, '/root/repo')
from lbgcmb_b200 import snr, _ffi, fisher
L = np.arange(50., 5001.); c = 1e-7 / L; v = 1e-16 / L; nm = np.ones_like(L)
for lmax in (1e5, 1200.0, 1200.0, 2400.0, 1e5):
    t = time.perf_counter(); r = snr.snr(L, c, v, nm, lmax=lmax); dt = time.perf_counter() - t
    print('snr lmax', lmax, '%.3f ms' % (dt * 1e3), 'kernel ms', _ffi.context().last_kernel_ms())
