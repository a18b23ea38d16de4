"""Peer-memory exchange (csrc/xchg.cu, dist.exchange) against the NCCL route on the same payloads, under torchrun on N >= 2 GPUs:
  timeout 300 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 scripts/xchg_check.py
Checks: curves identical, sums equal to the rank-order sum (and to NCCL's to 1e-15), ragged and empty shards, 300 back-to-back
exchanges with rank-dependent delays (the double-buffered windows), then times both routes at the bench's payload."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    import torch
    import torch.distributed as dist
    rank, world, local = (int(os.environ[k]) for k in ('RANK', 'WORLD_SIZE', 'LOCAL_RANK'))
    torch.cuda.set_device(local)
    dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from lbgcmb_b200 import dist as ld
    rng = np.random.default_rng(100 + rank)

    def both(sums, curves, n):
        a = ld.exchange(sums, curves, n)
        assert ld.last_exchange_backend == 'p2p', (ld.last_exchange_backend, ld._p2p['why'])
        keep = ld._p2p['ctx']
        ld._p2p['ctx'] = None                           # the NCCL route on the same payload
        b = ld.exchange(sums, curves, n)
        assert ld.last_exchange_backend == 'nccl'
        ld._p2p['ctx'] = keep
        return a, b

    bad = 0
    for n in (4951, 37, world - 1 if world > 1 else 1, 1, 2 * world + 1):
        mine = ld.shard_indices(n)
        full = np.arange(10 * n, dtype=np.float64).reshape(10, n) * 1.25 + 3.0
        sums = rng.normal(size=(4, 9))
        (s1, c1, ms1), (s2, c2, ms2) = both(sums, full[:, mine], n)
        allsums = [None] * world
        dist.all_gather_object(allsums, sums)
        want = allsums[0].copy()
        for r in range(1, world):
            want += allsums[r]
        ok = np.array_equal(c1, full) and np.array_equal(c2, full) and np.array_equal(s1, want) and np.allclose(s2, want, rtol=1e-14, atol=1e-15)
        bad += not ok
        if rank == 0:
            print('n = %d: curves identical %s, sums rank-order %s, vs NCCL %.1e' % (n, np.array_equal(c1, full), np.array_equal(s1, want),
                                                                                     float(np.max(np.abs(s1 - s2)))))
    # back-to-back exchanges with skewed arrival: every result checked
    n = 4951
    mine = ld.shard_indices(n)
    for k in range(300):
        if (k + rank) % 7 == 0:
            time.sleep(0.002 * ((rank + k) % 3))
        full = np.full((10, n), float(k)) + np.arange(n)[None, :]
        sums = np.full((4, 9), float(rank + 1) * (k + 1))
        s, c, _ = ld.exchange(sums, full[:, mine], n)
        ok = np.array_equal(c, full) and np.array_equal(s, np.full((4, 9), (k + 1) * world * (world + 1) / 2.))
        bad += not ok
    # timing at the bench's payload (36 sums + 10 curves of the rank's shard)
    def timed(reps=200):
        dev, wall = [], []
        for _ in range(reps):
            dist.barrier()
            t0 = time.perf_counter()
            _, _, ms = ld.exchange(sums, full[:, mine], n)
            wall.append((time.perf_counter() - t0) * 1e6)
            dev.append(ms * 1e3)
        return float(np.median(dev)), float(np.median(wall))
    p2p = timed()
    keep = ld._p2p['ctx']
    ld._p2p['ctx'] = None
    nccl = timed()
    ld._p2p['ctx'] = keep
    tot = torch.tensor([float(bad)], device='cuda')
    dist.all_reduce(tot)
    if rank == 0:
        print(json.dumps({'world': world, 'mismatches': int(tot.item()), 'p2p_device_us': p2p[0], 'p2p_wall_us': p2p[1],
                          'nccl_device_us': nccl[0], 'nccl_wall_us': nccl[1]}))
    dist.barrier()
    dist.destroy_process_group()


if __name__ == '__main__':
    main()
