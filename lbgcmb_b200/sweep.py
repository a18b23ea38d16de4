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


_FINE = {}


def _fine_grid(zmin, zmax):
    """np.arange(zmin, zmax, 0.001): the normalisation grid of sliced_pz.py:6 / pz2nbar.py:8, kept per range."""
    g = _FINE.get((zmin, zmax))
    if g is None:
        if len(_FINE) > 16:
            _FINE.clear()
        g = _FINE[(zmin, zmax)] = np.arange(zmin, zmax, 0.001)
        g.setflags(write=False)
    return g


_DZ = {}


def _grid_dz(zs):
    """The spacing sliced_pz.py:22-24 divides by: the 6-decimal rounded step of the (asserted uniform) grid."""
    key = (zs.size, float(zs[0]), float(zs[-1]))
    dz = _DZ.get(key)
    if dz is None:
        steps = set(np.around(np.ediff1d(zs), 6))
        assert len(steps) == 1
        if len(_DZ) > 16:
            _DZ.clear()
        dz = _DZ[key] = float(list(steps)[0])
    return dz


def forecast_spectra(Pk_interps, Llls, samples, zmin=0.01, zmax=10.0, want_ckk=True):
    """(C_kk[n_l] or None, C_gg[n, n_l], C_kg[n, n_l], N_gg[n]) for samples = [{'pz', 'bz', 'nbar'}, ...] on a common z range:
    what lensing.Ckk, Cgg.Cgg(..., zeff=False), Ckg.Ckg(..., zeff=False) and Cgg.Ngg return per sample -- the body of
    interloper.get_allCls (interloper.py:23-40) for a batch, in ONE library call (lbg_forecast_spectra).  The host evaluates each
    callable once per grid (the C_gg nodes, the C_kg nodes, and the dz = 0.001 normalisation grid shared by sliced_pz.py:6-9 and
    pz2nbar.py:6-17); the p(z) renormalisation, the prefactors, the quadrature weights and the sums over z run on the device."""
    from lbgcmb_b200.Cgg import _cgg_nodes, _each
    from lbgcmb_b200.Ckg import _ckg_nodes
    from lbgcmb_b200.lensing import _ckk_nodes
    from lbgcmb_b200.params import get_params
    from lbgcmb_b200.pz2nbar import nbar_convert
    par = get_params()
    engine, nonlinear = engine_of(Pk_interps, 'nlinear')
    Ld = _ffi.f64(Llls)
    zg, chig, wg, hcg = _cgg_nodes(float(zmin), float(zmax))
    zk, chik, wk, wkap = _ckg_nodes(float(zmin), float(zmax))
    n_s = len(samples)
    grids = [zg, zk]
    if want_ckk:
        zc, chic, wc = _ckk_nodes(float(par['zscatter']))
        grids.append(zc)
    else:
        chic = wc = np.zeros(0)
    # Order of the host work (the GPU builds the P(k,z) tables meanwhile): the linear table goes up and its sigma^2 sums start;
    # the callables are evaluated on the two Limber grids; the table build is enqueued; the callables are evaluated on the
    # normalisation grid (the larger half of the host work); then the one library call of the stage.
    ctx = engine._ctx()
    pgg, bgg, pkg, bkg = np.empty((n_s, zg.size)), np.empty((n_s, zg.size)), np.empty((n_s, zk.size)), np.empty((n_s, zk.size))
    norm, ngg = np.empty(n_s), np.empty(n_s)
    for i, smp in enumerate(samples):
        pz, bz = smp['pz'], smp['bz']
        pgg[i], pkg[i] = pz(zg), pz(zk)
        bgg[i], bkg[i] = _each(bz, zg), _each(bz, zk)
    slots = engine.build_tables(grids, nonlinear=nonlinear, wait=False)
    fine = _fine_grid(float(zmin), float(zmax))
    for i, smp in enumerate(samples):
        norm[i] = np.trapezoid(smp['pz'](fine), fine)           # sliced_pz.py:6-9 == pz2nbar.pz_slice (same grid, same rule)
        ngg[i] = 1. / (nbar_convert(smp['nbar'], unit='str') * norm[i]) if smp.get('nbar') is not None else np.nan   # Cgg.py:85-95
    cgg, ckg = np.empty((n_s, Ld.size)), np.empty((n_s, Ld.size))
    ckk = np.empty(Ld.size) if want_ckk else None
    dp = _ffi._dp
    chic, wc = _ffi.f64(chic), _ffi.f64(wc)
    for attempt in (0, 1):
        slot_g, slot_k = slots[0], slots[1]
        slot_c = slots[2] if want_ckk else 0
        _ffi.check(_ffi.lib().lbg_forecast_spectra(
            ctx.handle, Ld.size, dp(Ld), float(zmin), float(zmax), float(par['h_100']),
            zg.size, dp(zg), dp(chig), dp(hcg), dp(wg), _grid_dz(zg), int(slot_g),
            zk.size, dp(zk), dp(chik), dp(wkap), dp(wk), _grid_dz(zk), int(slot_k),
            chic.size, dp(chic), dp(wc), int(slot_c),
            n_s, dp(pgg), dp(bgg), dp(pkg), dp(bkg), dp(norm), dp(cgg), dp(ckg), dp(ckk) if want_ckk else None))
        if engine.finish_tables():
            break
        slots = engine.build_tables(grids, nonlinear=nonlinear)     # a redshift took the reference's exception path: explicit rows, again
    return ckk, cgg, ckg, ngg


def limber_grid(Pk_interps, Llls, samples, zmin=0.01, zmax=10.0):
    """C_gg[n, n_l], C_kg[n, n_l], N_gg[n] for samples = [{'pz', 'bz', 'nbar'}, ...] on a common z range (forecast_spectra
    without C_kk)."""
    return forecast_spectra(Pk_interps, Llls, samples, zmin, zmax, want_ckk=False)[1:]


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
        def spectra_fn(smp):
            return forecast_spectra(Pk_interps, Ld, smp, zmin, zmax)
    if grid_fn is None:
        grid_fn = forecast_grid
    lo, hi = _dist.shard_slice(len(noise_configs))
    mine = list(noise_configs[lo:hi])
    groups = []
    if n0_async and mine:
        # this rank's noise levels grouped by beam: levels of one beam share every table read of the N0 integrand
        # (Nkk.n0_levels, csrc/n0.cu n0_rows5m_kernel); one call per beam, on the second context while the spectra are built
        from lbgcmb_b200.Nkk import n0_levels_begin
        by_beam = {}
        for i, cfg in enumerate(mine):
            by_beam.setdefault(float(cfg[0]), []).append(i)
        groups = list(by_beam.items())

        class _PerLevel(object):            # a beam the multi-level kernel cannot serve (beyond its folded-noise bound): level by level
            def __init__(self, beam, idx):
                self.beam, self.idx = beam, idx

            def result(self):
                from lbgcmb_b200.Nkk import n0_terms
                return (np.array([n0_terms(lensCl_interps, nolensCl_interps, Llls, terms, self.beam, mine[i][1], mine[i][2], lane=1)[0]
                                  for i in self.idx]),)

        def start(gi):
            beam, idx = groups[gi]
            try:
                return n0_levels_begin(lensCl_interps, nolensCl_interps, Llls, beam, [mine[i][1] for i in idx],
                                       [int(bool(mine[i][2])) for i in idx], lane=1)
            except ValueError:
                return _PerLevel(beam, idx)
        if tuple(terms) != ('TT', 'TE', 'EE', 'EB'):
            raise ValueError('survey_sweep evaluates the four estimators TT, TE, EE, EB')
        call = start(0)
    ckk, cgg, ckg, ngg = spectra_fn(samples)
    xc = spectra_fn(xsamples)[1:] if xsamples is not None else (None, None, None)
    if n0_async:
        nkk = np.empty((hi - lo, Ld.size))
        for gi in range(len(groups)):              # one call in flight per context: collect group gi, start group gi + 1
            res = call.result()[0]
            nkk[groups[gi][1]] = res
            if gi + 1 < len(groups):
                call = start(gi + 1)
    else:
        nkk = np.array([n0_fn(cfg) for cfg in mine]).reshape(hi - lo, Ld.size)
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
                     timing=None, overlap=True):
    """ONE forecast (one experiment; one galaxy sample or a list of them) with the multipole grid sharded over the ranks -- the
    layout north_star names: every rank evaluates N0, C_kk, C_gg, C_kg and the per-multipole covariance / Fisher terms on its own
    multipoles (dist.shard_indices: L[rank::world], because the N0 cost per L falls with L; N0 per L and Limber per l are
    independent), and the only exchange inside the path is ONE all-reduce of the partial sums (per sample: S/N^2, the 2 x 2 Fisher
    matrix of (sigma8, b1), the bias vector, N_gg) over NCCL.  The curves are then gathered (one all_gather) so that every rank
    returns the full arrays, as snr.py / interloper.py produce them.  On a GPU the N0 curve is started first on the library's second
    context (Nkk.n0_terms_async: enqueued, not waited for) and the spectra are built while it runs; `overlap=False` makes the
    calls strictly one after another (per-stage device timing).

    noise_config = (thetab, DeltaT, iterative); sample = {'pz', 'bz', 'nbar', 'fid_b1', 'LllMax'} or a list of such; xsample: the
    interloper-contaminated twin(s) (bias vector).  Returns {'snr', 'Fisher', 'interim', 'dtheta', 'ngg', 'nkk', 'ckk', 'cgg',
    'ckg'}; with a list of samples every per-sample entry gains a leading sample axis.
    `n0_fn`, `spectra_fn`, `grid_fn` replace the device calls in the CPU (gloo) test of the sharding logic; `timing`, a dict,
    receives this rank's wall seconds per phase and the device ms of the exchange."""
    import time
    t0 = time.perf_counter()
    single = isinstance(sample, dict)
    samples = [sample] if single else list(sample)
    xsamples = None if xsample is None else ([xsample] if isinstance(xsample, dict) else list(xsample))
    n_b = len(samples)
    if xsamples is not None and len(xsamples) != n_b:
        raise ValueError('xsample must pair up with sample')
    Ld = np.asarray(Llls, dtype=np.float64)
    nm = np.asarray(nmodes, dtype=np.float64)
    mine = _dist.shard_indices(Ld.size)
    Lr = Ld[mine]
    have = mine.size > 0
    if spectra_fn is None:
        def spectra_fn(Ls, smp):
            return forecast_spectra(Pk_interps, Ls, smp, zmin, zmax)
    if grid_fn is None:
        grid_fn = forecast_grid
    smp = samples if xsamples is None else samples + xsamples
    if have:
        if n0_fn is None:
            from lbgcmb_b200.Nkk import n0_begin
            call = n0_begin(lensCl_interps, nolensCl_interps, np.asarray(Llls)[mine], terms, noise_config[0], noise_config[1],
                            noise_config[2], lane=1 if overlap else 0)
            if not overlap:
                call.result()
            ckk, cgg, ckg, ngg = spectra_fn(Lr, smp)
            t1 = time.perf_counter()
            nkk = call.result()[0]
        else:
            nkk = np.asarray(n0_fn(np.asarray(Llls)[mine]), dtype=np.float64)
            ckk, cgg, ckg, ngg = spectra_fn(Lr, smp)
            t1 = time.perf_counter()
        t2 = time.perf_counter()
        x = {} if xsamples is None else {'xcgg': cgg[n_b:], 'xckg': ckg[n_b:], 'xngg': ngg[n_b:]}
        part = grid_fn(Lr, nm[mine], ckk, cgg[:n_b], ckg[:n_b], ngg[:n_b], nkk[None, :], fsky, fid_sig8,
                       [s_['fid_b1'] for s_ in samples], [s_['LllMax'] for s_ in samples], **x)[0]         # [n_b, 8]
        sums = np.empty((n_b, 9))
        sums[:, 0] = part[:, 0] ** 2.                                 # S/N adds in quadrature over the shards
        sums[:, 1:7] = part[:, 1:7]
        sums[:, 7] = ngg[:n_b]
        sums[:, 8] = 1.0
        curves = np.concatenate([nkk[None, :], ckk[None, :], cgg[:n_b], ckg[:n_b]])
    else:                                                             # more ranks than multipoles
        t1 = t2 = time.perf_counter()
        sums = np.zeros((n_b, 9))
        curves = np.zeros((2 + 2 * n_b, 0))
    t3 = time.perf_counter()
    sums, curves, xms = _dist.exchange(sums, curves, Ld.size)         # THE exchange step of the path: one all-reduce, one all-gather
    t4 = time.perf_counter()
    F = sums[:, 1:5].reshape(n_b, 2, 2)
    interim = sums[:, 5:7]
    dtheta = np.zeros((n_b, 2))
    if xsamples is not None:
        ok = np.linalg.det(F) != 0                                    # stacked calls: LAPACK per matrix, as one call per sample gives
        if ok.any():
            dtheta[ok] = np.linalg.solve(F[ok], interim[ok][:, :, None])[:, :, 0]
    if timing is not None:
        timing.update(spectra_s=t1 - t0, n0_wait_s=t2 - t1, grid_s=t3 - t2, collectives_s=t4 - t3, exchange_device_ms=xms,
                      n_L_mine=int(mine.size))
    res = {'snr': np.sqrt(sums[:, 0]), 'Fisher': F, 'interim': interim, 'dtheta': dtheta, 'ngg': sums[:, 7] / sums[:, 8],
           'nkk': curves[0], 'ckk': curves[1], 'cgg': curves[2:2 + n_b], 'ckg': curves[2 + n_b:2 + 2 * n_b]}
    if single:
        res = {k: (v if k in ('nkk', 'ckk') else v[0]) for k, v in res.items()}
        res['snr'] = float(res['snr'])
    return res
