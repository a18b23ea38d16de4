"""Survey-design sweep (BASELINE.json configs[4]) on N GPUs: beam 1-7' x noise 1-30 uK' x 64 LBG n(z) configurations.

  python scripts/sweep_bench.py [--n-noise 16] [--grid A|B]                       (one GPU)
  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/sweep_bench.py ...

64 n(z) configurations = 4 bands x 4 widths {0.15, 0.24, 0.35, 0.5} x 4 nbar scalings {0.25, 0.5, 1, 2}.
Noise levels: the first --n-noise of the 7 x 30 (beam, DeltaT) grid in a fixed seed-0 permutation (--full: all 210, beam-major), iterative-EB
for DeltaT <= 3.  Levels of one beam go through the multi-level N0 kernel (Nkk.n0_levels).
Prints one JSON line on rank 0 (wall-clock of the whole sweep, max over ranks through the final gather)."""
import argparse
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--n-noise', type=int, default=16)
    ap.add_argument('--grid', default='B', choices=['A', 'B'])
    ap.add_argument('--full', action='store_true', help='the stated size of configs[4]: all 7 x 30 = 210 (beam, noise) levels, beam-major')
    a = ap.parse_args()
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', '0'), ('WORLD_SIZE', '1'), ('LOCAL_RANK', '0')))
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    import bench
    from lbgcmb_b200 import pmh, sweep, _ffi
    from lbgcmb_b200.prep_Llls import prep_Llls
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    from lbgcmb_b200.synthetic import DROPOUTS, gaussian_pz, linz_bz
    li, ni = cl_interps_from_tables(*bench.make_inputs())
    kk, pp, _ = bench.forecast_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp, device=local))
    _, Ls, nmodes = prep_Llls(60, 50., 5000., log10=(a.grid == 'A'))
    samples, xsamples = [], []
    for name, (z0, s0, nbar) in DROPOUTS.items():
        for sig in (0.15, 0.24, 0.35, 0.5):
            for f in (0.25, 0.5, 1.0, 2.0):
                pz = gaussian_pz(z0, sig)
                pzi = gaussian_pz(max(0.3 * (1. + z0) - 1., 0.1), 0.1)
                pzz = (lambda p, q: (lambda z: 0.94 * p(z) + 0.06 * q(z)))(pz, pzi)
                samples.append({'pz': pz, 'bz': linz_bz, 'nbar': f * nbar, 'fid_b1': 1. + z0, 'LllMax': bench.LLLMAX[name]})
                xsamples.append({'pz': pzz, 'bz': linz_bz, 'nbar': 1.06 * f * nbar})
    noise = [(tb, dT, dT <= 3.0) for tb in range(1, 8) for dT in range(1, 31)]
    perm = np.random.default_rng(0).permutation(len(noise))
    noise = sorted(noise) if a.full else [noise[i] for i in perm[:a.n_noise]]      # beam-major: a rank's block holds whole beams
    ctx, ctx1 = _ffi.context(local), _ffi.context(local, 1)
    sweep.survey_sweep(li, ni, Pk, Ls[:8], nmodes[:8], noise[:world], samples[:2], bench.FSKY, xsamples=xsamples[:2])     # warm-up
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ctx.kernel_totals(reset=True)
    ctx1.kernel_totals(reset=True)
    t0 = time.perf_counter()
    res = sweep.survey_sweep(li, ni, Pk, Ls, nmodes, noise, samples, bench.FSKY, xsamples=xsamples)
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    kms, launches = ctx.kernel_totals()
    kms1, launches1 = ctx1.kernel_totals()
    assert np.all(np.isfinite(res['out'])) and np.all(res['out'][:, :, 0] > 0)
    if rank == 0:
        n_design = len(noise) * len(samples)
        print(json.dumps({'workload': 'survey sweep: %d noise levels x %d n(z) configurations on %d multipoles' % (len(noise), len(samples), Ls.size),
                          'n_gpus': world, 'wall_s': dt, 'designs_per_s': n_design / dt, 'n0_L_modes_per_s': len(noise) * Ls.size / dt,
                          'rank0_kernel_ms': kms + kms1, 'rank0_n0_kernel_ms': kms1, 'rank0_launches': launches + launches1,
                          'best_snr_design': [int(x) for x in np.unravel_index(np.argmax(res['out'][:, :, 0]), res['out'].shape[:2])],
                          'snr_range': [float(res['out'][:, :, 0].min()), float(res['out'][:, :, 0].max())]}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
