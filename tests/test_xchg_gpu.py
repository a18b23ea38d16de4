"""GPU test of the peer-memory exchange (csrc/xchg.cu) through the C ABI: two PROCESSES, each with its own context on cuda:0 (the
windows are mapped with CUDA IPC exactly as between two GPUs; on one GPU the two kernels are time-sliced, which also exercises the
wait-for-a-late-peer path).  Ragged and empty shards, repeated exchanges (double-buffered windows), argument errors.
The N-GPU check against the NCCL route is scripts/xchg_check.py (profiles/r2p_xchg_check_n*.log)."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, qs, out):
    sys.path.insert(0, ROOT)
    import ctypes as C
    from lbgcmb_b200 import _ffi
    lib = _ffi.lib()
    ctx = _ffi.context(0)
    handle = (C.c_ubyte * 64)()
    _ffi.check(lib.lbg_xchg_create(ctx.handle, rank, world, 4096, handle))
    for r in range(world):
        if r != rank:
            qs[r].put((rank, bytes(handle)))
    got = {rank: bytes(handle)}
    while len(got) < world:
        r, h = qs[rank].get(timeout=60)
        got[r] = h
    blob = (C.c_ubyte * (64 * world)).from_buffer_copy(b''.join(got[r] for r in range(world)))
    _ffi.check(lib.lbg_xchg_open(ctx.handle, blob))
    ok = True
    for k, n in enumerate([11, 1, 0, 257, 11, 11, 64]):
        mine = np.arange(rank, n, world)
        full = np.arange(3 * n, dtype=np.float64).reshape(3, n) * 0.5 + k
        sums = np.array([1.0 + rank + k, 0.1 * (rank + 1)])
        red, curves, ms = np.empty(2), np.empty((3, n)), C.c_double(0.0)
        _ffi.check(lib.lbg_xchg_sums_curves(ctx.handle, 2, _ffi._dp(sums), 3, n, _ffi._dp(np.ascontiguousarray(full[:, mine])), _ffi._dp(red),
                                            _ffi._dp(curves), C.byref(ms)))
        want = np.array([1.0 + 0 + k, 0.1])
        for r in range(1, world):
            want = want + np.array([1.0 + r + k, 0.1 * (r + 1)])
        ok = ok and np.array_equal(curves, full) and np.array_equal(red, want) and ms.value >= 0.0
    block = np.full(5, float(rank))
    allb = np.empty((world, 5))
    _ffi.check(lib.lbg_xchg_allgather(ctx.handle, 5, _ffi._dp(block), _ffi._dp(allb), None))
    ok = ok and np.array_equal(allb, np.repeat(np.arange(world, dtype=np.float64)[:, None], 5, axis=1))
    too_big = lib.lbg_xchg_allgather(ctx.handle, 1 << 20, _ffi._dp(np.zeros(1 << 20)), _ffi._dp(np.zeros(2 << 20)), None)
    ok = ok and too_big == -1
    _ffi.check(lib.lbg_xchg_destroy(ctx.handle))
    out.put((rank, bool(ok)))


def test_peer_memory_exchange_two_processes(ctx):
    import torch.multiprocessing as mp
    mpc = mp.get_context('spawn')
    world = 2
    qs = [mpc.Queue() for _ in range(world)]
    out = mpc.Queue()
    procs = [mpc.Process(target=_worker, args=(r, world, qs, out)) for r in range(world)]
    for p in procs:
        p.start()
    res = dict(out.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res == {0: True, 1: True}


def test_exchange_state_errors(ctx):
    import ctypes as C
    from lbgcmb_b200 import _ffi
    lib = _ffi.lib()
    h = (C.c_ubyte * 64)()
    assert lib.lbg_xchg_allgather(ctx.handle, 4, _ffi._dp(np.zeros(4)), _ffi._dp(np.zeros(8)), None) == -3        # no window yet
    assert lib.lbg_xchg_create(ctx.handle, 0, 1, 64, h) == -1 and lib.lbg_xchg_create(ctx.handle, 2, 2, 64, h) == -1
    assert lib.lbg_xchg_create(ctx.handle, 0, 2, 64, h) == 0
    assert lib.lbg_xchg_allgather(ctx.handle, 4, _ffi._dp(np.zeros(4)), _ffi._dp(np.zeros(8)), None) == -3        # peer window not opened
    assert lib.lbg_xchg_destroy(ctx.handle) == 0
