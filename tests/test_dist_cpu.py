"""CPU, world_size 2 over gloo: the sharding helpers and the Fisher all-reduce (the one exchange step of the path).
The per-shard arithmetic is replaced by a numpy stand-in here; on the GPU box the same code runs lbg_fisher_accum."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _numpy_accum(cov, dD, w, dDelta=None):
    icov = np.linalg.inv(cov) if len(w) else np.zeros_like(cov)
    F = np.einsum('l,lpi,lij,lqj->pq', w, dD, icov, dD)
    b = None if dDelta is None else np.einsum('l,lpi,lij,lj->p', w, dD, icov, dDelta)
    return F, b


def _problem(n=37, ns=3, npar=2):
    rng = np.random.default_rng(4)
    A = rng.normal(size=(n, ns, ns))
    cov = A @ np.transpose(A, (0, 2, 1)) + ns * np.eye(ns)
    return cov, rng.normal(size=(n, npar, ns)), rng.integers(1, 50, n).astype(float), rng.normal(size=(n, ns))


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from lbgcmb_b200 import dist as ld
    cov, dD, w, dl = _problem()
    F, b = ld.fisher_accum_sharded(cov, dD, w, dDelta=dl, accum=_numpy_accum)
    F2, b2 = ld.fisher_accum_sharded(cov, dD, w, accum=_numpy_accum)
    lo, hi = ld.shard_slice(len(w))
    curve = ld.allgather_concat(w[lo:hi] * 2.0)
    # cyclic shards: the one-round-trip exchange of a sharded forecast
    n = 11
    mine = ld.shard_indices(n)
    full = np.arange(3 * n, dtype=np.float64).reshape(3, n) + 0.5
    sums, curves, ms = ld.exchange(np.array([1.0 + rank, 10.0]), full[:, mine], n)
    ok = np.array_equal(sums, [3.0, 20.0]) and np.array_equal(curves, full) and ms is None
    q.put((rank, F, b, F2, b2 is None, (lo, hi), curve, ok))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_slices_partition_everything():
    from lbgcmb_b200.dist import shard_slice
    for n in (0, 1, 7, 60, 4951):
        for world in (1, 2, 3, 8):
            cuts = [shard_slice(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(cuts[r][1] == cuts[r + 1][0] for r in range(world - 1))
            sizes = [hi - lo for lo, hi in cuts]
            assert max(sizes) - min(sizes) <= 1


def test_single_process_is_identity():
    from lbgcmb_b200 import dist as ld
    cov, dD, w, dl = _problem()
    F, b = ld.fisher_accum_sharded(cov, dD, w, dDelta=dl, accum=_numpy_accum)
    Fr, br = _numpy_accum(cov, dD, w, dl)
    assert np.array_equal(F, Fr) and np.array_equal(b, br)
    assert ld.rank_world() == (0, 1)


def test_fisher_allreduce_world2_gloo():
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted([q.get(timeout=120) for _ in procs], key=lambda t: t[0])
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    cov, dD, w, dl = _problem()
    Fr, br = _numpy_accum(cov, dD, w, dl)
    for rank, F, b, F2, b2none, cut, curve, ok in res:
        assert ok
        assert np.allclose(F, Fr, rtol=1e-13) and np.allclose(b, br, rtol=1e-12)
        assert np.allclose(F2, Fr, rtol=1e-13) and b2none
        assert np.array_equal(curve, w * 2.0)
    assert res[0][5] == (0, 19) and res[1][5] == (19, 37)
    assert np.array_equal(res[0][1], res[1][1])            # every rank holds the same reduced matrix
