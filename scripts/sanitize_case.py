"""One small call of every kernel family, for compute-sanitizer (memcheck / racecheck):
  compute-sanitizer --tool memcheck python scripts/sanitize_case.py
Sizes are tiny (the sanitizer slows kernels 10-100x) but walk the same code paths as the bench: N0 rows5 / rows3 + special rows + second pass,
the modi N0, the Limber kernel in both batch templates with out-of-range nodes, halofit table slots, S/N, covariance, Fisher, the
design grid, mqw / cadence Fisher, E-H and growth tables, the multi-level N0 kernel, the fused forecast spectra, spectra kept on the device
and the field Fisher sum fed from them, the S/N design grid.  (The peer-memory exchange needs two processes: tests/test_xchg_gpu.py.)"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from lbgcmb_b200 import Nkk, eisenstein_hu, interloper, pmh, snr, sweep
from lbgcmb_b200.fisher import fisher_accum, gaussian_cov
from lbgcmb_b200.modi.cosmology import growth_tables
from lbgcmb_b200.prep_camb import cl_interps_from_tables

li, ni = cl_interps_from_tables(*bench.make_inputs())
Ls = np.array([50, 2999, 3001, 5000])
nkk, terms = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True)
nkk2 = Nkk.n0_terms_async(li, ni, Ls[:2], ['TT', 'EB'], 7.0, 30.0, False).result()[0]
for fold in (2, False):        # the generic per-node-mask kernel on the folded and on the full phi grid
    assert np.allclose(Nkk.n0_terms(li, ni, Ls[:2], ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=fold)[0], nkk[:2], rtol=1e-11)
Nkk.n0_terms(li, ni, Ls[:2], ['TT', 'BB'], 9.0, 30.0, False)      # analytic-noise variant (wide beam) + BB estimator
kk, pp, bands = bench.forecast_inputs()
Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp))
Lf = np.array([50., 300., 1500., 4000.])
rows = bench.forecast_stage((interloper, snr), Pk, bands[:2], Lf, np.ones_like(Lf), 1e-8 * np.ones_like(Lf))
smp = [{'pz': b['pz'], 'bz': b['bz'], 'nbar': b['nbar'], 'fid_b1': b['fid_b1'], 'LllMax': 2000.} for b in bands] * 3   # 12 entries: NB = 8 + ragged tail
cgg, ckg, ngg = sweep.limber_grid(Pk, Lf, smp)
out = sweep.forecast_grid(Lf, np.ones_like(Lf), 1e-7 / Lf, cgg, ckg, ngg, np.stack([nkk, 2 * nkk]), 0.4, 0.819, [s['fid_b1'] for s in smp], [2000.] * len(smp))
rng = np.random.default_rng(0)
A = rng.normal(size=(7, 5, 5))
cov = A @ np.transpose(A, (0, 2, 1)) + 5 * np.eye(5)
F, b = fisher_accum(cov, rng.normal(size=(7, 3, 5)), np.ones(7), dDelta=rng.normal(size=(7, 5)))
tot = np.tile(np.array([[2., .5], [.5, 1.]]), (4, 1, 1))
gaussian_cov(Lf, tot, [(0, 0), (0, 1), (1, 1)], 0.4)
k = np.logspace(-4, 1, 300)
eisenstein_hu.linear_power(k, 0.292, 0.0468, 0.708, 0.69, 0.965, 0.819)
growth_tables([0.25, 0.3])
print('ok', nkk, rows.shape, out.shape, F.shape)
# round 2 additions
lev, _ = Nkk.n0_levels(li, ni, Ls[:3], 1.4, [1.0, 7.0, 3.0, 12.0], [True, False, True, False])          # multi-level kernel, groups of 3 + 1
assert np.allclose(lev[0], nkk[:3], rtol=1e-11)
sweep.forecast_spectra(Pk, Lf, smp[:4])                                                                  # device-built weights + three Limber sums
from lbgcmb_b200 import fisher, tomography as tm
dndz = lambda z: z ** 2. * np.exp(-(z / 0.5))
tr = [{'name': 'a', 'edges': np.linspace(0.4, 1.0, 4), 'dndz': dndz, 'bz': lambda z: 1.1 + 0.4 * z, 'nbar': 3e6}]
r = tm.tomographic_fisher(tr, Lf, np.ones_like(Lf), 0.4, 2e-8 * np.ones_like(Lf), cosmo_params=('sig8', 'mnu'))      # spectra sets -> lbg_field_fisher_sets
assert np.all(np.isfinite(r['F'])) and fisher.spectra_set_get(0, 0, 1, Lf.size).shape == (1, 4)
g = snr.snr_grid(Lf, ckg[0], 1e-7 / Lf, cgg[0], np.stack([nkk, 2 * nkk]), [0.4, 0.2], np.zeros((2, 4)), [1e-9, 1e-8, 1e-7], np.ones_like(Lf))
assert g.shape == (2, 2, 3) and np.all(g > 0)
print('sanitize_case: every kernel family ran')
