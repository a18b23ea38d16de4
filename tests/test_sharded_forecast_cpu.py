"""CPU, gloo: sweep.sharded_forecast -- one forecast with the multipole grid split over ranks and ONE all-reduce of the partial
S/N^2 / Fisher / bias sums.  Device stages replaced by numpy stand-ins whose sums over L are exactly additive,
so world sizes 1, 2 and 3 (and 5 > number of multipoles in the small case) must agree to rounding."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LS = np.arange(50., 61.)
SAMPLE = {'fid_b1': 3.5, 'LllMax': 58.0}


def _n0(Ls):
    return 1e-8 * (1. + np.asarray(Ls, dtype=np.float64) / 100.)


def _spectra(Ls, smp):
    n = len(smp)
    return 1e-7 / Ls, np.outer(np.arange(1, n + 1), 1e-5 / Ls), np.outer(np.arange(1, n + 1), 1e-6 / Ls ** 0.5), 1e-7 * np.arange(1., n + 1.)


def _grid(Ld, nmodes, ckk, cgg, ckg, ngg, nkk, fsky, fid_sig8, fid_b1, LllMax, xcgg=None, xckg=None, xngg=None):
    out = np.zeros((1, 1, 8))
    keep = Ld < LllMax[0]
    r = ckg[0] ** 2. / ((ckk + nkk[0]) * (cgg[0] + ngg[0]))
    out[0, 0, 0] = np.sqrt(np.sum(nmodes * r))
    out[0, 0, 1] = np.sum((nmodes * Ld)[keep])
    out[0, 0, 2] = out[0, 0, 3] = np.sum((nmodes * np.sqrt(Ld))[keep])
    out[0, 0, 4] = 3. * np.sum((nmodes * Ld ** 1.5)[keep])
    if xcgg is not None:
        out[0, 0, 5] = np.sum((nmodes * (xcgg[0] - cgg[0]))[keep]) * 1e5
        out[0, 0, 6] = np.sum((nmodes * (xckg[0] - ckg[0]))[keep]) * 1e6
    return out


def _run(Ls=LS):
    from lbgcmb_b200 import sweep
    return sweep.sharded_forecast(None, None, None, Ls, np.ones_like(Ls), (1.4, 1.0, True), SAMPLE, 0.4, xsample=dict(SAMPLE),
                                  n0_fn=_n0, spectra_fn=_spectra, grid_fn=_grid)


def _worker(rank, world, port, q, n_l):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    res = _run(LS[:n_l])
    q.put((rank, {k: np.asarray(v) for k, v in res.items()}))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_forecast_worlds_agree():
    for n_l, worlds in ((LS.size, (2, 3)), (3, (5,))):
        ref = _run(LS[:n_l])
        assert ref['nkk'].shape == (n_l,) and ref['Fisher'].shape == (2, 2) and ref['snr'] > 0
        assert np.array_equal(ref['nkk'], _n0(LS[:n_l]))
        for world in worlds:
            ctx = mp.get_context('spawn')
            q = ctx.Queue()
            port = 29640 + world + n_l
            procs = [ctx.Process(target=_worker, args=(r, world, port, q, n_l)) for r in range(world)]
            for p in procs:
                p.start()
            got = [q.get(timeout=180) for _ in procs]
            for p in procs:
                p.join(timeout=60)
                assert p.exitcode == 0
            for rank, res in got:
                for key in ('nkk', 'ckk', 'cgg', 'ckg'):
                    assert np.array_equal(res[key], ref[key]), (world, rank, key)
                for key in ('snr', 'Fisher', 'interim', 'dtheta', 'ngg'):
                    assert np.allclose(res[key], ref[key], rtol=1e-13, atol=0), (world, rank, key)
