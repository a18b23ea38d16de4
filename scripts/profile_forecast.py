"""Host-side phase breakdown of one forecast step at a given shard size (what bounds the strong scaling of bench.py --gpus N):
  python scripts/profile_forecast.py [stride]      stride = 8 -> the 619-multipole shard one rank of an 8-GPU run evaluates
Wall-clock microseconds per phase, median of 20 steps, N0 overlapped on the second context as in the bench."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import Nkk, _ffi, pmh, sweep
from lbgcmb_b200.prep_camb import cl_interps_from_tables

stride = int(sys.argv[1]) if len(sys.argv) > 1 else 1
li, ni = cl_interps_from_tables(*bench.make_inputs())
kk, pp, bands = bench.forecast_inputs()
samples, xsamples = bench.band_samples(bands)
smp = samples + xsamples
T = pmh.PkTable(kk, pp)
Pk = pmh.get_PkInterps(T)
Ls = np.arange(50, 5001)[::stride]
Lf = Ls.astype(np.float64)
ctx1 = _ffi.context(None, 1)
rec = {}


def tick(name, t0):
    rec.setdefault(name, []).append((time.perf_counter() - t0) * 1e6)


for rep in range(25):
    T.engine.invalidate_device()
    t_all = time.perf_counter()
    t0 = time.perf_counter(); Nkk.upload_cls(ctx1, li, ni); tick('upload_cls (stack + compare)', t0)
    t0 = time.perf_counter(); call = Nkk.n0_begin(li, ni, Ls, bench.TERMS, 1.4, 1.0, True, lane=1); tick('n0_begin (incl. upload_cls again)', t0)
    t0 = time.perf_counter(); ckk, cgg, ckg, ngg = sweep.forecast_spectra(Pk, Lf, smp); tick('forecast_spectra (upload, callables, tables built meanwhile, 1 library call)', t0)
    t0 = time.perf_counter(); nkk = call.result()[0]; tick('n0 wait', t0)
    t0 = time.perf_counter()
    sweep.forecast_grid(Lf, np.ones_like(Lf), ckk, cgg[:4], ckg[:4], ngg[:4], nkk[None, :], 0.4, 0.819, [s['fid_b1'] for s in samples],
                        [s['LllMax'] for s in samples], xcgg=cgg[4:], xckg=ckg[4:], xngg=ngg[4:])
    tick('forecast_grid', t0)
    tick('TOTAL', t_all)
out = {k: float(np.median(v[5:])) for k, v in rec.items()}
print(json.dumps({'n_L': int(Ls.size), 'stride': stride, 'median_us': out}, indent=1))
