"""HALOFIT spectral parameters of the forecast's three z grids written to a .npy file: run once per library build
(LBGCMB_LIB=...) to compare kernel versions bit for bit.  usage: python scripts/halofit_spec_dump.py out.npy"""
import os
import sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import bench
from lbgcmb_b200 import pmh
from lbgcmb_b200.Cgg import _cgg_nodes
from lbgcmb_b200.Ckg import _ckg_nodes
from lbgcmb_b200.lensing import _ckk_nodes
kk, pp, bands = bench.forecast_inputs()
T = pmh.PkTable(kk, pp)
out = []
for zs in (_cgg_nodes(0.01, 10.0)[0], _ckg_nodes(0.01, 10.0)[0], _ckk_nodes(1100.)[0]):
    zs = np.asarray(zs)
    zs = zs[zs < 4.0]
    D2 = np.array([float(T.engine.Dgrow(z=z)) ** 2. for z in zs])
    out.append(T.engine.halofit_specs(zs, D2))
out = np.concatenate(out)
np.save(sys.argv[1], out)
print(out.shape, out[:2], out[-1])
