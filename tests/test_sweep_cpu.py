"""CPU, gloo: the sharding / gather logic of sweep.survey_sweep (noise levels split over ranks, results concatenated in
rank order).  The three device stages are replaced by deterministic numpy stand-ins; on the GPU box the same driver runs
lbg_n0_ho02, lbg_limber and lbg_forecast_grid (tests/test_sweep_gpu.py)."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
NOISE = [(1.0 + 0.5 * i, 1.0 + i, i % 2 == 0) for i in range(5)]
SAMPLES = [{'fid_b1': 3.0 + g, 'LllMax': 1000. + 100 * g} for g in range(3)]
LS = np.arange(50., 58.)


def _n0(cfg):
    return (cfg[0] + 10. * cfg[1] + 100. * cfg[2]) * LS


def _spectra(smp):
    n = len(smp)
    return LS * 1e-3, np.outer(np.arange(1, n + 1), LS), np.outer(np.arange(1, n + 1), LS ** 0.5), np.arange(1., n + 1.)


def _grid(Ld, nmodes, ckk, cgg, ckg, ngg, nkk, fsky, fid_sig8, fid_b1, LllMax, xcgg=None, xckg=None, xngg=None):
    out = np.zeros((nkk.shape[0], cgg.shape[0], 8))
    out[:, :, 0] = nkk.sum(axis=1)[:, None] + cgg.sum(axis=1)[None, :]
    out[:, :, 1] = np.asarray(fid_b1)[None, :] * fsky
    return out


def _run():
    from lbgcmb_b200 import sweep
    return sweep.survey_sweep(None, None, None, LS, np.ones_like(LS), NOISE, SAMPLES, 0.4,
                              n0_fn=_n0, spectra_fn=_spectra, grid_fn=_grid)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    res = _run()
    q.put((rank, res['out'], res['nkk']))
    dist.barrier()
    dist.destroy_process_group()


def test_sweep_single_process():
    res = _run()
    assert res['out'].shape == (5, 3, 8) and res['nkk'].shape == (5, 8)
    assert np.array_equal(res['nkk'][3], _n0(NOISE[3]))


def test_sweep_world2_and_world3_match_single_process():
    ref = _run()
    for world in (2, 3):
        ctx = mp.get_context('spawn')
        q = ctx.Queue()
        port = 31500 + os.getpid() % 2000 + world
        procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
        for p in procs:
            p.start()
        res = [q.get(timeout=180) for _ in procs]
        for p in procs:
            p.join(timeout=60)
            assert p.exitcode == 0
        for rank, out, nkk in res:
            assert np.array_equal(out, ref['out']) and np.array_equal(nkk, ref['nkk'])
