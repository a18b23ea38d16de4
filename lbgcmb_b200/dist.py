"""Multi-GPU plumbing: one process per GPU (torchrun), multipole / configuration sharding, and the only exchange step of
the path -- the sum of the per-shard Fisher matrix and bias vector.

`torch.distributed` is used for the rendezvous and the collective (NCCL over NVLink on GPUs; gloo in the CPU tests);
the per-shard arithmetic is lbg_fisher_accum on each rank's own GPU.  Units (N0 per L, Limber per (pair, l)) are
independent, so nothing else communicates."""
import numpy as np


def _dist():
    import torch.distributed as dist
    return dist if (dist.is_available() and dist.is_initialized()) else None


def rank_world():
    d = _dist()
    return (d.get_rank(), d.get_world_size()) if d else (0, 1)


def shard_slice(n, rank=None, world=None):
    """Contiguous block [lo, hi) of n units for this rank (sizes differ by at most one; cost per L is uniform)."""
    if rank is None or world is None:
        rank, world = rank_world()
    base, extra = divmod(n, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_indices(n, rank=None, world=None):
    """Cyclic shard: units rank, rank + world, ... of n.  For N0 the cost per L falls with L (more of the (l1, phi) grid is masked
    by the lmax cuts and skipped), so contiguous blocks would leave the low-L rank with the longest job."""
    if rank is None or world is None:
        rank, world = rank_world()
    return np.arange(rank, n, world)


def allgather_cyclic(local, n):
    """Inverse of shard_indices for any number of equally sharded float64 curves: `local` is [n_curves, n_mine] on this rank;
    returns [n_curves, n] on every rank, one collective (NCCL all_gather of a padded block; gloo on CPU)."""
    d = _dist()
    local = np.atleast_2d(np.ascontiguousarray(local, dtype=np.float64))
    if d is None or d.get_world_size() == 1:
        return local.copy()
    import torch
    world = d.get_world_size()
    width = (n + world - 1) // world
    pad = np.zeros((local.shape[0], width))
    pad[:, :local.shape[1]] = local
    dev = torch.device('cuda', torch.cuda.current_device()) if d.get_backend() == 'nccl' else torch.device('cpu')
    mine = torch.from_numpy(pad).to(dev)
    parts = torch.empty((world * pad.shape[0], width), dtype=torch.float64, device=dev)      # ranks concatenated along dim 0
    d.all_gather_into_tensor(parts, mine)
    parts = parts.cpu().numpy().reshape(world, pad.shape[0], width)
    out = np.empty((local.shape[0], n))
    for r in range(world):
        m = len(range(r, n, world))
        out[:, r::world] = parts[r, :, :m]
    return out


_xbuf = {}
last_exchange_ms = None
last_exchange_backend = None        # 'p2p' (peer-memory kernel, csrc/xchg.cu), 'nccl', 'gloo' or None (single rank)
_P2P_CAP = 1 << 16                  # doubles per rank and exchange (512 KB); larger payloads take the NCCL route
_p2p = {'tried': False, 'ctx': None, 'why': None}


def _p2p_context():
    """The library context whose peer-memory windows serve `exchange` on NCCL process groups, or None (then `_p2p['why']` says why).
    Set up once per process, collectively: every rank creates its window (lbg_xchg_create), the 64-byte IPC handles go round with
    all_gather_object, every rank opens its peers' windows (lbg_xchg_open); the route is used only if every rank succeeded.
    LBGCMB_XCHG=nccl keeps the exchange on NCCL."""
    if _p2p['tried']:
        return _p2p['ctx']
    _p2p['tried'] = True
    import ctypes as C
    import os
    d = _dist()
    if os.environ.get('LBGCMB_XCHG', 'p2p').lower() == 'nccl':
        _p2p['why'] = 'LBGCMB_XCHG=nccl'
        return None
    from lbgcmb_b200 import _ffi
    rank, world = d.get_rank(), d.get_world_size()
    ctx, handle, why = None, (C.c_ubyte * 64)(), None
    try:
        ctx = _ffi.context()
        rc = _ffi.lib().lbg_xchg_create(ctx.handle, rank, world, _P2P_CAP, handle)
        if rc != 0:
            why = _ffi.lib().lbg_last_error().decode('utf-8', 'replace')
    except Exception as e:                                  # noqa: BLE001 (any failure here means "use NCCL", the reason is kept)
        why = repr(e)
    got = [None] * world
    d.all_gather_object(got, (why, bytes(handle)))
    if any(g[0] for g in got):
        _p2p['why'] = next(g[0] for g in got if g[0])
    else:
        blob = (C.c_ubyte * (64 * world)).from_buffer_copy(b''.join(g[1] for g in got))
        rc = _ffi.lib().lbg_xchg_open(ctx.handle, blob)
        why = None if rc == 0 else _ffi.lib().lbg_last_error().decode('utf-8', 'replace')
        got = [None] * world
        d.all_gather_object(got, why)
        if any(got):
            _p2p['why'] = next(g for g in got if g)
        else:
            _p2p['ctx'] = ctx
    if _p2p['ctx'] is None and ctx is not None:
        _ffi.lib().lbg_xchg_destroy(ctx.handle)
    return _p2p['ctx']


def exchange(sums, curves, n):
    """The exchange step of a sharded forecast in one round trip: `sums` (any shape) is all-reduced (SUM) and `curves`
    [n_curves, n_mine] (this rank's cyclic shard of n units, shard_indices) is all-gathered back into [n_curves, n].

    On GPUs (NCCL process group) the exchange runs over NVLink peer memory when the windows could be set up (_p2p_context): ONE
    kernel per rank stores [sums | curves] into every peer's window and waits for the peers' blocks (csrc/xchg.cu); the sums are then
    added on the host in rank order -- the all-reduce, with a summation order that does not depend on the collective library.
    Otherwise (LBGCMB_XCHG=nccl, set-up failed, payload beyond the window) both payloads travel in ONE pinned host block -> one device
    tensor, an NCCL all_reduce and an all_gather_into_tensor run on slices of it, and one copy brings the result back.  The device
    time of the exchange is returned as the third value (None off-GPU).  gloo / no process group: CPU tensors / identity."""
    global last_exchange_ms, last_exchange_backend
    d = _dist()
    sums = np.ascontiguousarray(sums, dtype=np.float64)
    curves = np.atleast_2d(np.ascontiguousarray(curves, dtype=np.float64))
    if d is None or d.get_world_size() == 1:
        last_exchange_ms = None
        last_exchange_backend = None
        return sums.copy(), curves.copy(), None
    import torch
    world = d.get_world_size()
    width = (n + world - 1) // world
    nc, ns = curves.shape[0], sums.size
    nccl = d.get_backend() == 'nccl'
    if nccl and ns + nc * width <= _P2P_CAP:
        ctx = _p2p_context()
        if ctx is not None:
            import ctypes as C
            from lbgcmb_b200 import _ffi
            if curves.shape[1] != len(range(d.get_rank(), n, world)):
                raise ValueError('curves must hold this rank\'s cyclic shard: %d of %d units' % (len(range(d.get_rank(), n, world)), n))
            red, out, ms = np.empty(ns), np.empty((nc, n)), C.c_double(0.0)
            flat = np.ascontiguousarray(curves)
            _ffi.check(_ffi.lib().lbg_xchg_sums_curves(ctx.handle, ns, _ffi._dp(sums.ravel()), nc, n, _ffi._dp(flat), _ffi._dp(red),
                                                       _ffi._dp(out), C.byref(ms)))
            last_exchange_ms, last_exchange_backend = ms.value, 'p2p'
            return red.reshape(sums.shape), out, ms.value
    key = (ns, nc, width, world, nccl)
    buf = _xbuf.get(key)
    if buf is None:
        n_in, n_out = ns + nc * width, ns + world * nc * width
        if nccl:
            dev = torch.device('cuda', torch.cuda.current_device())
            buf = (torch.empty(n_in, dtype=torch.float64).pin_memory(), torch.empty(n_out, dtype=torch.float64).pin_memory(),
                   torch.empty(n_in, dtype=torch.float64, device=dev), torch.empty(n_out, dtype=torch.float64, device=dev),
                   torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
        else:
            buf = (torch.empty(n_in, dtype=torch.float64), torch.empty(n_out, dtype=torch.float64), None, None, None, None)
        if len(_xbuf) >= 4:                              # a driver alternates between at most a few shapes; keep their pinned blocks
            _xbuf.clear()
        _xbuf[key] = buf
    h_in, h_out, d_in, d_out, e0, e1 = buf
    hin = h_in.numpy()
    hin[:ns] = sums.ravel()
    pad = hin[ns:].reshape(nc, width)
    pad[:, :curves.shape[1]] = curves
    pad[:, curves.shape[1]:] = 0.0
    if nccl:
        d_in.copy_(h_in, non_blocking=True)
        e0.record()
        d_out[:ns].copy_(d_in[:ns])
        d.all_reduce(d_out[:ns], op=d.ReduceOp.SUM)
        d.all_gather_into_tensor(d_out[ns:], d_in[ns:])
        e1.record()
        h_out.copy_(d_out, non_blocking=True)
        torch.cuda.current_stream().synchronize()
        ms = e0.elapsed_time(e1)
    else:
        h_out[:ns] = h_in[:ns]
        d.all_reduce(h_out[:ns], op=d.ReduceOp.SUM)
        d.all_gather_into_tensor(h_out[ns:], h_in[ns:].clone())
        ms = None
    hout = h_out.numpy()
    parts = hout[ns:].reshape(world, nc, width)
    out = np.empty((nc, n))
    for r in range(world):
        m = len(range(r, n, world))
        out[:, r::world] = parts[r, :, :m]
    last_exchange_ms, last_exchange_backend = ms, ('nccl' if nccl else 'gloo')
    return hout[:ns].reshape(sums.shape).copy(), out, ms


def allreduce_sum(*arrays):
    """Sum float64 numpy arrays over all ranks in one collective (packed); identity without a process group."""
    d = _dist()
    if d is None or d.get_world_size() == 1:
        return [np.array(a, dtype=np.float64) for a in arrays]
    import torch
    flat = np.concatenate([np.asarray(a, dtype=np.float64).ravel() for a in arrays])
    dev = torch.device('cuda', torch.cuda.current_device()) if d.get_backend() == 'nccl' else torch.device('cpu')
    t = torch.from_numpy(flat).to(dev)
    d.all_reduce(t, op=d.ReduceOp.SUM)
    flat = t.cpu().numpy()
    out, o = [], 0
    for a in arrays:
        a = np.asarray(a)
        out.append(flat[o:o + a.size].reshape(a.shape).copy())
        o += a.size
    return out


def allgather_concat(local):
    """Concatenate each rank's 1-d float64 slice in rank order (per-L N0 / C_l curves when every rank needs the whole)."""
    d = _dist()
    local = np.ascontiguousarray(local, dtype=np.float64)
    if d is None or d.get_world_size() == 1:
        return local.copy()
    parts = [None] * d.get_world_size()
    d.all_gather_object(parts, local)
    return np.concatenate(parts)


def fisher_accum_sharded(cov, dD, mode_weight, dDelta=None, accum=None):
    """interloper.py:223-253 with the multipoles sharded over ranks: every rank passes the FULL arrays, evaluates its own
    block on its GPU and receives the all-reduced (F, b).  `accum` replaces the device call in the CPU (gloo) tests."""
    if accum is None:
        from lbgcmb_b200.fisher import fisher_accum as accum
    lo, hi = shard_slice(len(mode_weight))
    F, b = accum(np.asarray(cov)[lo:hi], np.asarray(dD)[lo:hi], np.asarray(mode_weight)[lo:hi],
                 dDelta=None if dDelta is None else np.asarray(dDelta)[lo:hi])
    if b is None:
        return allreduce_sum(F)[0], None
    F, b = allreduce_sum(F, b)
    return F, b
