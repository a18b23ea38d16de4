"""BASELINE.json configs[3] at its stated size with its stated parameters: kappa + 6 tracers x 20 redshift bins (dz = 0.1) on the 4951
multipoles L = 50..5000, Fisher matrix over (sigma8, sum m_nu) + one bias amplitude per tracer (8 parameters, 121 fields, 241 spectra).

  python scripts/tomography_fisher_bench.py            (one GPU; prints one JSON line)
  torchrun --nproc-per-node N scripts/tomography_fisher_bench.py     (multipoles sharded cyclically, one all-reduce of the 8 x 8 matrix)

Five cosmologies (fiducial, sigma8 +-, m_nu +-) are built on the device (E-H linear table with massive neutrinos, growth tables, HALOFIT
rows), each projected to 241 spectra; the Fisher sum is one lbg_field_fisher launch.  Tracers are synthetic stand-ins as in
scripts/tomography_bench.py."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

TRACERS = {   # name: (z0, sigma, nbar [deg^-2], b0)
    'BGS': (0.3, 0.15, 1400., 1.34), 'LRG': (0.75, 0.15, 600., 2.1), 'ELG': (1.0, 0.25, 2400., 1.2),
    'QSO': (1.7, 0.6, 170., 2.3), 'LBGu': (2.96, 0.24, 2181., 3.96), 'LBGg': (3.8, 0.3, 5409., 4.8)}
N_BIN, DZ = 20, 0.1


def main():
    import bench
    from lbgcmb_b200 import Nkk, _ffi, dist, tomography as tm
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    world = int(os.environ.get('WORLD_SIZE', '1'))
    if world > 1:
        import torch
        import torch.distributed as td
        torch.cuda.set_device(int(os.environ.get('LOCAL_RANK', '0')))
        td.init_process_group('nccl')
    rank = dist.rank_world()[0]
    li, ni = cl_interps_from_tables(*bench.make_inputs())
    Ls = np.arange(50, 5001)
    Lf = Ls.astype(np.float64)
    tracers = []
    for name, (z0, sig, nbar, b0) in TRACERS.items():
        lo0 = max(z0 - N_BIN * DZ / 2., 0.05)
        edges = lo0 + DZ * np.arange(N_BIN + 1)
        pz = (lambda z0_, s_: (lambda z: np.exp(-0.5 * ((np.asarray(z) - z0_) / s_) ** 2.)))(z0, sig)
        zz = np.arange(0.0, 10.0, 0.001)
        frac = np.array([np.trapezoid(pz(zz[(zz >= a) & (zz <= b)]), dx=0.001) for a, b in zip(edges[:-1], edges[1:])]) / np.trapezoid(pz(zz), dx=0.001)
        tracers.append({'name': name, 'edges': edges, 'dndz': pz, 'bz': (lambda b_, z_: (lambda z: b_ * (1. + z) / (1. + z_)))(b0, z0),
                        'nbar': np.maximum(nbar * frac, 1e-3) * (180. / np.pi) ** 2.})
    ctx = _ffi.context()
    idx = dist.shard_indices(Ls.size)
    nkk = np.zeros(Ls.size)
    nkk[idx] = Nkk.Nkk(li, ni, Ls[idx], ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, iterative=True, pickleit=False)
    small = [dict(t, edges=t['edges'][:3], nbar=t['nbar'][:2]) for t in tracers[:2]]
    tm.tomographic_fisher(small, Lf[:64], np.ones(64), bench.FSKY, nkk[:64] + 1e-7, cosmo_params=('sig8',), sharded=world > 1)      # warm-up (incl. the first NCCL all-reduce)
    tf = time.perf_counter()
    tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), bench.FSKY, nkk, cosmo_params=('sig8', 'mnu'), sharded=world > 1)   # first call at this size: buffers grow
    first_ms = (time.perf_counter() - tf) * 1e3
    ctx.kernel_totals(reset=True)
    t0 = time.perf_counter()
    r = tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), bench.FSKY, nkk, cosmo_params=('sig8', 'mnu'), sharded=world > 1)
    t1 = time.perf_counter()
    kms, launches = ctx.kernel_totals()
    ff_ms = ctx.last_kernel_ms()[0]
    # a caller that loops over noise curves keeps the geometry (distances, weight vectors) between calls
    geo = {}
    tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), bench.FSKY, nkk, cosmo_params=('sig8', 'mnu'), sharded=world > 1, geometry=geo)
    t2 = time.perf_counter()
    r2 = tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), bench.FSKY, 1.5 * nkk, cosmo_params=('sig8', 'mnu'), sharded=world > 1, geometry=geo,
                               return_spectra=False)
    t3 = time.perf_counter()
    assert np.all(np.diag(r2['F']) < np.diag(r['F']))                  # more lensing noise, less information
    fs8 = tm.fsigma8_forecast(tm.tomographic_fisher(tracers, Lf, np.ones_like(Lf), bench.FSKY, nkk, cosmo_params=('sig8', 'M'), sharded=world > 1,
                                                    geometry={}), [0.5, 1.0, 2.0, 3.0])
    F = r['F']
    assert np.all(np.isfinite(F)) and np.all(np.linalg.eigvalsh(F) > 0)
    if rank == 0:
        print(json.dumps({'workload': 'kappa + %d tracers x %d bins (dz = %.1f) on %d multipoles: %d fields, %d spectra x %d cosmologies, '
                                      'Fisher over %s' % (len(TRACERS), N_BIN, DZ, Ls.size, r['fields'], 2 * r['fields'] - 1,
                                                          r['n_cosmologies'], r['params']),
                          'n_gpus': world, 'wall_ms': (t1 - t0) * 1e3, 'first_call_wall_ms': first_ms, 'wall_ms_geometry_kept_no_readback': (t3 - t2) * 1e3,
                          'fsigma8_rel_sigma_sig8_M': dict(zip(['z=0.5', 'z=1', 'z=2', 'z=3'], (fs8['sigma'] / fs8['fsigma8']).tolist())), 'kernel_ms_this_rank': kms, 'launches_this_rank': launches,
                          'field_fisher_kernel_ms': ff_ms, 'sigma': dict(zip(r['params'], r['sigma'].tolist())),
                          'F_diag': np.diag(F).tolist()}), flush=True)
    if world > 1:
        td.destroy_process_group()


if __name__ == '__main__':
    main()
