"""Survey-design sweep (BASELINE.json configs[4]): N0 + S/N + (sigma8, b1) Fisher over beam x noise x n(z) configurations,
sharded over the GPUs of one box.

The reference has no such driver: it loops `Nkk` over experiments (snr.py:108-116, bolometers.py) and re-runs its scripts
per dropout band.  Here the same per-design arithmetic -- Nkk.Nkk per (thetab, DeltaT, iterative); Cgg.Cgg / Ckg.Ckg
(zeff=False) / lensing.Ckk per n(z); snr.snr on the Ckg.var_Ckg variance; interloper.get_ClsCov + the Fisher / bias sums of
interloper.py:200-260 -- runs as
  * one N0 launch per noise level (csrc/n0.cu; the dominant cost, independent per level -> the unit that is sharded),
  * ONE batched Limber launch per spectrum type for all n(z) configurations (csrc/limber.cu, batch axis),
  * ONE launch for the whole noise x n(z) grid of S/N + Fisher problems (csrc/sweep.cu, lbg_forecast_grid).
Multi-GPU: rank r takes a contiguous block of noise levels (dist.shard_slice); no data-path collective;
the [n_noise, n_cfg, 8] result blocks are concatenated in rank order at the end (dist.allgather_concat)."""
import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200 import dist as _dist
from lbgcmb_b200.Cgg import Ngg, cgg_nodes, cgg_weights
from lbgcmb_b200.Ckg import ckg_nodes, ckg_weights
from lbgcmb_b200.pmh import engine_of

OUT_FIELDS = ('snr', 'F_s8s8', 'F_s8b1', 'F_b1s8', 'F_b1b1', 'bias_s8', 'bias_b1')


def limber_grid(Pk_interps, Llls, samples, zmin=0.01, zmax=10.0):
    """C_gg[n, n_l], C_kg[n, n_l], N_gg[n] for samples = [{'pz', 'bz', 'nbar'}, ...] on a common z range: what
    Cgg.Cgg(..., zeff=False), Ckg.Ckg(..., zeff=False) and Cgg.Ngg return per sample, from two batched launches."""
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = np.asarray(Llls, dtype=np.float64)
    zs, chis, w = cgg_nodes(zmin, zmax)
    wg = np.stack([cgg_weights(zs, chis, w, zmin, zmax, s['pz'], s['bz']) for s in samples])
    cgg = engine.limber(Ld, zs, chis, wg, nonlinear=nonlinear)
    zs, chis, w = ckg_nodes(zmin, zmax)
    wk = np.stack([ckg_weights(zs, chis, w, zmin, zmax, s['pz'], s['bz']) for s in samples])
    ckg = engine.limber(Ld, zs, chis, wk, nonlinear=nonlinear)
    ngg = np.array([Ngg(Ld[:1], zmin, zmax, s['pz'], s['nbar'])[0] for s in samples])
    return cgg, ckg, ngg


def forecast_grid(Llls, nmodes, ckk, cgg, ckg, ngg, nkk, fsky, fid_sig8, fid_b1, LllMax, snr_lmax=None,
                  xcgg=None, xckg=None, xngg=None, device=None):
    """out[n_noise, n_cfg, 8] = (S/N, F_s8s8, F_s8b1, F_b1s8, F_b1b1, bias_s8, bias_b1, 0) for every (N0 curve, n(z) config).

    Per design this equals snr.snr(Llls, ckg, var_Ckg, nmodes, lmax=snr_lmax) and interloper.fisher_bias(...)['Fisher'],
    ['interim']; the contaminated spectra are optional (no bias vector without them)."""
    ctx = _ffi.context(device)
    L, nm, ckk = _ffi.f64(Llls), _ffi.f64(nmodes), _ffi.f64(ckk)
    cgg, ckg, nkk = _ffi.f64(np.atleast_2d(cgg)), _ffi.f64(np.atleast_2d(ckg)), _ffi.f64(np.atleast_2d(nkk))
    n_cfg, n_ell = cgg.shape
    if ckg.shape != cgg.shape or nkk.shape[1] != n_ell or L.shape != (n_ell,) or nm.shape != (n_ell,) or ckk.shape != (n_ell,):
        raise ValueError('spectra / multipole arrays disagree in shape')
    cfg = np.zeros((n_cfg, 8))
    cfg[:, 0] = ngg
    cfg[:, 1] = ngg if xngg is None else xngg
    cfg[:, 2] = fid_b1
    cfg[:, 3] = LllMax
    cfg[:, 4] = LllMax if snr_lmax is None else snr_lmax
    bias = xcgg is not None
    if bias:
        xcgg, xckg = _ffi.f64(np.atleast_2d(xcgg)), _ffi.f64(np.atleast_2d(xckg))
        if xcgg.shape != cgg.shape or xckg.shape != cgg.shape:
            raise ValueError('contaminated spectra disagree in shape')
    out = np.empty((nkk.shape[0], n_cfg, 8))
    _ffi.check(_ffi.lib().lbg_forecast_grid(ctx.handle, n_ell, _ffi._dp(L), _ffi._dp(nm), _ffi._dp(ckk), n_cfg, _ffi._dp(cgg),
                                            _ffi._dp(ckg), _ffi._dp(xcgg) if bias else None, _ffi._dp(xckg) if bias else None,
                                            _ffi._dp(_ffi.f64(cfg)), nkk.shape[0], _ffi._dp(nkk), float(fsky), float(fid_sig8),
                                            _ffi._dp(out)))
    return out


def survey_sweep(lensCl_interps, nolensCl_interps, Pk_interps, Llls, nmodes, noise_configs, samples, fsky,
                 fid_sig8=0.819, xsamples=None, zmin=0.01, zmax=10.0, terms=('TT', 'TE', 'EE', 'EB'),
                 n0_fn=None, spectra_fn=None, grid_fn=None):
    """Full grid: noise_configs = [(thetab, DeltaT, iterative), ...] x samples = [{'pz','bz','nbar','fid_b1','LllMax'}, ...]
    (xsamples: the interloper-contaminated twins, same length, for the bias vector).

    Returns {'out': [n_noise, n_cfg, 8] (OUT_FIELDS), 'nkk': [n_noise, n_l], 'ckk', 'cgg', 'ckg', 'ngg'} on every rank.
    Under torch.distributed each rank evaluates its own block of noise levels.  `n0_fn`, `spectra_fn`, `grid_fn` replace the
    device calls in the CPU (gloo) tests of the sharding logic."""
    Ld = np.asarray(Llls, dtype=np.float64)
    n0_async = n0_fn is None
    if spectra_fn is None:
        from lbgcmb_b200.lensing import Ckk

        def spectra_fn(smp):
            return (Ckk(Pk_interps, Ld),) + limber_grid(Pk_interps, Ld, smp, zmin, zmax)
    if grid_fn is None:
        grid_fn = forecast_grid
    lo, hi = _dist.shard_slice(len(noise_configs))
    futs = None
    if n0_async:                                 # this rank's N0 curves run on the second context while the spectra are built
        from lbgcmb_b200.Nkk import n0_terms_async
        futs = [n0_terms_async(lensCl_interps, nolensCl_interps, Llls, terms, cfg[0], cfg[1], cfg[2], reuse_tables=True)
                for cfg in noise_configs[lo:hi]]
    ckk, cgg, ckg, ngg = spectra_fn(samples)
    xc = spectra_fn(xsamples)[1:] if xsamples is not None else (None, None, None)
    if futs is not None:
        nkk = np.array([f.result()[0] for f in futs]).reshape(hi - lo, Ld.size)
    else:
        nkk = np.array([n0_fn(cfg) for cfg in noise_configs[lo:hi]]).reshape(hi - lo, Ld.size)
    if hi > lo:
        out = grid_fn(Ld, nmodes, ckk, cgg, ckg, ngg, nkk, fsky, fid_sig8, [s['fid_b1'] for s in samples],
                      [s['LllMax'] for s in samples], xcgg=xc[0], xckg=xc[1], xngg=xc[2])
    else:
        out = np.zeros((0, len(samples), 8))
    out = _dist.allgather_concat(out.ravel()).reshape(len(noise_configs), len(samples), 8)
    nkk = _dist.allgather_concat(nkk.ravel()).reshape(len(noise_configs), Ld.size)
    return {'out': out, 'nkk': nkk, 'ckk': ckk, 'cgg': cgg, 'ckg': ckg, 'ngg': ngg}


def sharded_forecast(lensCl_interps, nolensCl_interps, Pk_interps, Llls, nmodes, noise_config, sample, fsky, fid_sig8=0.819,
                     xsample=None, zmin=0.01, zmax=10.0, terms=('TT', 'TE', 'EE', 'EB'), n0_fn=None, spectra_fn=None, grid_fn=None,
                     timing=None):
    """ONE forecast (one experiment, one galaxy sample) with the multipole grid sharded over the ranks -- the layout north_star
    names: every rank evaluates N0, C_kk, C_gg, C_kg and the per-multipole covariance / Fisher terms on its own multipoles
    (dist.shard_indices: L[rank::world], because the N0 cost per L falls with L; N0 per L and Limber per l are independent), and
    the only exchange inside the path is ONE all-reduce of the partial sums (S/N^2, the 2 x 2 Fisher matrix of (sigma8, b1), the
    bias vector, N_gg): 10 doubles over NCCL.  The curves are then gathered (one all_gather) so that every rank returns the full
    arrays, as snr.py / interloper.py produce them.  On a GPU the N0 curve runs on the library's second context
    (Nkk.n0_terms_async) while the spectra are built.

    noise_config = (thetab, DeltaT, iterative); sample = {'pz', 'bz', 'nbar', 'fid_b1', 'LllMax'}; xsample: the interloper-
    contaminated twin (bias vector).  Returns {'snr', 'Fisher', 'interim', 'dtheta', 'nkk', 'ckk', 'cgg', 'ckg', 'ngg'}.
    `n0_fn`, `spectra_fn`, `grid_fn` replace the device calls in the CPU (gloo) test of the sharding logic; `timing`, a dict,
    receives this rank's wall seconds per phase."""
    import time
    t0 = time.perf_counter()
    Ld = np.asarray(Llls, dtype=np.float64)
    nm = np.asarray(nmodes, dtype=np.float64)
    mine = _dist.shard_indices(Ld.size)
    Lr = Ld[mine]
    have = mine.size > 0
    if spectra_fn is None:
        from lbgcmb_b200.lensing import Ckk

        def spectra_fn(Ls, smp):
            return (Ckk(Pk_interps, Ls),) + limber_grid(Pk_interps, Ls, smp, zmin, zmax)
    if grid_fn is None:
        grid_fn = forecast_grid
    smp = [sample] if xsample is None else [sample, xsample]
    if have:
        if n0_fn is None:
            from lbgcmb_b200.Nkk import n0_terms_async
            fut = n0_terms_async(lensCl_interps, nolensCl_interps, np.asarray(Llls)[mine], terms, noise_config[0], noise_config[1],
                                 noise_config[2])
            ckk, cgg, ckg, ngg = spectra_fn(Lr, smp)
            t1 = time.perf_counter()
            nkk = fut.result()[0]
        else:
            nkk = np.asarray(n0_fn(np.asarray(Llls)[mine]), dtype=np.float64)
            ckk, cgg, ckg, ngg = spectra_fn(Lr, smp)
            t1 = time.perf_counter()
        t2 = time.perf_counter()
        x = {} if xsample is None else {'xcgg': cgg[1:2], 'xckg': ckg[1:2], 'xngg': ngg[1:2]}
        part = grid_fn(Lr, nm[mine], ckk, cgg[:1], ckg[:1], ngg[:1], nkk[None, :], fsky, fid_sig8, [sample['fid_b1']],
                       [sample['LllMax']], **x)[0, 0]
        sums = np.concatenate([[part[0] ** 2.], part[1:7], ngg[:1], [1.0]])     # S/N adds in quadrature over the shards
        curves = np.stack([nkk, ckk, cgg[0], ckg[0]])
    else:                                                             # more ranks than multipoles
        t1 = t2 = time.perf_counter()
        sums = np.zeros(9)
        curves = np.zeros((4, 0))
    t3 = time.perf_counter()
    sums = _dist.allreduce_sum(sums)[0]                               # THE exchange step of the path
    curves = _dist.allgather_cyclic(curves, Ld.size)
    t4 = time.perf_counter()
    F = sums[1:5].reshape(2, 2)
    interim = sums[5:7]
    if timing is not None:
        timing.update(spectra_s=t1 - t0, n0_wait_s=t2 - t1, grid_s=t3 - t2, collectives_s=t4 - t3, n_L_mine=int(mine.size))
    return {'snr': float(np.sqrt(sums[0])), 'Fisher': F, 'interim': interim,
            'dtheta': np.linalg.solve(F, interim) if xsample is not None and np.linalg.det(F) != 0 else np.zeros(2),
            'ngg': sums[7] / sums[8], 'nkk': curves[0], 'ckk': curves[1], 'cgg': curves[2], 'ckg': curves[3]}
