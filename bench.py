#!/usr/bin/env python
"""bench.py -- headline benchmark of the LBGCMB forecast hot path on B200 (see DESIGN.md sections 5 and 6).

One step = ONE forecast: the Hu-Okamoto N0 (TT+TE+EE+EB, iterative EB; BASELINE.json configs[1]: thetab=1.4', DeltaT=1 uK',
every integer L in [50, 5000] -> 4951 L-modes) followed by the Limber C_kk/C_kg/C_gg + covariance + S/N + (sigma8, b1)
Fisher / interloper-bias stage for the u, g, r, i dropout samples on the same multipoles (configs[0], [2]).
metric = L-modes per second.  The step is `sweep.sharded_forecast`, the call a user makes.

N = 1: the whole multipole grid on one GPU.
N > 1 (torchrun, one rank per GPU): the SAME forecast with the multipole grid sharded over the ranks (rank r takes L[r::N]), ONE
NCCL all-reduce of the partial S/N^2 / Fisher / bias sums and one all-gather of the curves inside the timed region -- the split
north_star names; "scaling": "strong" (total work fixed).

  value      whole-job L-modes/s over the DEVICE time of the step: every kernel of the step (CUDA events on the library's stream,
             calls made one after another; max over ranks) + the collectives (CUDA events on torch's stream; min over ranks: the
             rank that arrives last measures the exchange without the wait for its peers)
  e2e        the same metric through the public call with HOST inputs (tables, callables) and host results: table comparison /
             upload, callable evaluation, H2D, kernels (N0 started first on the library's second context and overlapped with the
             Limber stage), D2H, collectives -- wall clock between barriers, max over ranks
  roofline   dominant kernel (n0_rows5_kernel): FP64 pipe; frac = 145 op x UNMASKED nodes (counted by the kernel) /
             kernel time / measured DFMA peak; frac_executed = executed flop per node from the committed ncu capture
             (profiles/*_n0_rows5_ncu.json, tied to the hash of csrc/n0.cu; `stale` when the source changed since)
  cpu_baseline  the reference's own CPU code (staged under oracle/_ref by oracle/stage_ref.py) or its numpy port, bounded sample

`--impl reference` times the CPU arm with all host cores (rank 0 only).
"""
import argparse
import glob
import hashlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

OPS_PER_NODE = 145          # frozen convention, SURVEY.md 8(d) / BASELINE.md section 4
FP64_NOMINAL_TFLOPS = 37.2  # 148 SM x 64 DFMA/clk x 2 x 1.965 GHz

# (thetab [arcmin], DeltaT [uK-arcmin], iterative): the headline is SWEEP[0]; the replica rate (rates.sweep_*) gives rank r SWEEP[r]
SWEEP = [(1.4, 1.0, True), (1.4, 7.0, False), (1.5, 12.0, False), (7.0, 30.0, False),
         (1.0, 1.0, True), (2.0, 2.0, True), (3.0, 5.0, False), (5.0, 20.0, False)]
TERMS = ['TT', 'TE', 'EE', 'EB']
METRIC = 'N0+Limber L-modes/sec (FP64)'
WORKLOAD = ('N0 Nkk TT/TE/EE/EB iterative-EB thetab=%.1f DeltaT=%.1f on every L=%d..%d (x 9999 l1 x 201 phi), then Limber '
            'Ckk/Ckg/Cgg (halofit P(k,z), full z integrals) + covariance + S/N + (sigma8,b1) Fisher/interloper bias for u,g,r,i dropouts on the same L')


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--lmin', type=int, default=50)
    ap.add_argument('--lmax', type=int, default=5000)
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--no-replica-rate', action='store_true', help='skip the per-rank replica sweep rate (rates.sweep_*) at N > 1')
    ap.add_argument('--cpu-sample', type=int, default=4, help='number of L values timed on the CPU arm inside the b200 line')
    ap.add_argument('--ref-port', action='store_true', help='--impl reference: time the numpy port even when the reference is staged')
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md clocks line)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(',')])
            except Exception:
                pass
            time.sleep(0.1)

    def summary(self):
        self.stop_flag = True
        self.join(timeout=6)
        sm, mx, reasons = [], 0.0, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = max(mx, float(r[1]))
                for name, v in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), r[3:7]):
                    if v.lower().startswith('active'):
                        reasons.add(name)
            except Exception:
                pass
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': mx or None,
                'reasons': sorted(reasons), 'samples': len(sm)}


def make_inputs():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    return cmb_cl_tables()


FSKY = 0.4
INT_FRAC = 0.06            # interloper.py:139
LLLMAX = {'u': 1200.0, 'g': 1600.0, 'r': 2000.0, 'i': 2400.0}   # Zel'dovich L cuts per sample (stand-ins for zeldovich_Lmax.Lcutmax)


def forecast_inputs():
    """Host-side inputs of the Limber / SNR / Fisher stage (BASELINE.json configs[0], [2]): P(k) table and, per dropout
    band, the clean and interloper-contaminated (p(z), b(z), nbar) triplets of interloper.py:128-171."""
    from lbgcmb_b200.synthetic import eh_like_plin, gaussian_pz, linz_bz, DROPOUTS
    k, p = eh_like_plin()
    bands = []
    for name, (z0, sig, nbar) in DROPOUTS.items():
        pz = gaussian_pz(z0, sig)
        zi = max(0.3 * (1. + z0) - 1., 0.1)
        pzi = gaussian_pz(zi, 0.1)
        pzz = (lambda a, b: (lambda z: (1. - INT_FRAC) * a(z) + INT_FRAC * b(z)))(pz, pzi)
        bands.append({'name': name, 'pz': pz, 'pzz': pzz, 'bz': linz_bz, 'nbar': nbar, 'nbar_wint': (1. + INT_FRAC) * nbar,
                      'fid_b1': float(linz_bz(z0)), 'LllMax': LLLMAX[name]})
    return k, p, bands


def band_samples(bands):
    """(samples, xsamples) in the form sweep.sharded_forecast takes: clean and interloper-contaminated twins per band."""
    samples = [{'pz': b['pz'], 'bz': b['bz'], 'nbar': b['nbar'], 'fid_b1': b['fid_b1'], 'LllMax': b['LllMax']} for b in bands]
    xsamples = [{'pz': b['pzz'], 'bz': b['bz'], 'nbar': b['nbar_wint']} for b in bands]
    return samples, xsamples


def forecast_stage(mods, Pk, bands, Ls, nmodes, nkk):
    """The forecast stage through the reference-shaped per-function calls: Limber C_kk / C_kg / C_gg (full z integrals, zmin = 0.01,
    zmax = 10 as interloper.py:148-149), Gaussian covariance, cumulative S/N of C_kg and the (sigma8, b1) Fisher matrix + interloper
    bias for every dropout band -- the body of interloper.py:168-260 and snr.py:118-124 through the same-named functions.
    `mods` = (interloper, snr) modules of the implementation under test (lbgcmb_b200 mirrors: the cross-check of the fused path;
    OracleForecast / ReferenceForecast: the CPU arms).  Returns one row per band: S/N, Fisher (4), dtheta (2)."""
    interloper, snr = mods
    spectra = []
    for b in bands:
        spectra.append((interloper.get_allCls(Pk, Ls, b['nbar'], FSKY, 0.01, 10.0, b['pz'], b['bz'], zeff=False, nkk=nkk),
                        interloper.get_allCls(Pk, Ls, b['nbar_wint'], FSKY, 0.01, 10.0, b['pzz'], b['bz'], zeff=False, nkk=nkk)))
    rows = []
    for b, ((Cls, Nls), (xCls, xNls)) in zip(bands, spectra):
        vkg = ((Cls['kk'] + Nls['kk']) * (Cls['gg'] + Nls['gg']) + Cls['kg'] ** 2.) / ((2. * Ls + 1.) * FSKY)   # Ckg.py:83-87
        sn = snr.snr(Ls, Cls['kg'], vkg, nmodes, lmax=b['LllMax']) / np.sqrt(FSKY)                                # snr.py:122-123
        fb = interloper.fisher_bias(Ls, nmodes, Cls, Nls, xCls, xNls, FSKY, 0.819, b['fid_b1'], b['LllMax'])
        rows.append(np.concatenate([[sn], fb['Fisher'].ravel(), fb['dtheta']]))
    return np.array(rows)


def rows_of(res):
    """The same rows from what sweep.sharded_forecast returns."""
    return np.concatenate([(res['snr'] / np.sqrt(FSKY))[:, None], res['Fisher'].reshape(-1, 4), res['dtheta']], axis=1)


class OracleForecast(object):
    """CPU legs only: the numpy restatements (oracle/limber.py, oracle/halofit.py, oracle/fisher.py) behind the call surface
    forecast_stage() uses.  P(k,z) rows are cached per redshift, the way the reference's camb interpolator is built once.
    `reference=True`: C_gg / C_kg / C_kk / N_gg come from the UNMODIFIED reference modules (Cgg.py, Ckg.py, cmb/lensing.py through
    oracle/refload.py: /root/reference in the build container, the byte-compiled copy under oracle/_ref/refpy.zip on the GPU box)."""

    def __init__(self, k, p, reference=False):
        from oracle import limber as ol, fisher as ofi, halofit as oh, modi_lensing as om
        from lbgcmb_b200.synthetic import PARAMS
        self.ol, self.ofi = ol, ofi
        self.bg = ol.Background(PARAMS)
        mc = om.ModiCosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=k, plin=p)
        self.Pk = oh.ModiPk(k, p, lambda z: float(mc.Dgrow(z=z)), mc.cosmo, PARAMS['sig_8']).interps()
        self.snr = ofi.snr
        self.fisher_bias = ofi.fisher_bias
        self.ref = None
        if reference:
            from oracle import refload
            self.ref = (refload.load('Cgg'), refload.load('Ckg'), refload.load('lensing'))

    def get_allCls(self, Pk, Llls, nbar, fsky, zmin, zmax, pz, bz, zeff=False, nkk=None):
        ol, L = self.ol, np.asarray(Llls, dtype=np.float64)
        if self.ref is not None:
            Cgg, Ckg, lensing = self.ref
            cgg = Cgg.Cgg(Pk, L, zmin, zmax, pz, bz, zeff=zeff, bz2=bz, survey_pz2=pz)      # interloper.py:33
            ngg = Cgg.Ngg(L, zmin, zmax, pz, nbar)
            ckg = Ckg.Ckg(Pk, L, zmin, zmax, pz, bz, zeff=zeff)
            ckk = lensing.Ckk(Pk, L)
        else:
            cgg = ol.Cgg(self.bg, Pk, L, zmin, zmax, pz, bz, zeff=zeff)
            ngg = ol.Ngg(L, zmin, zmax, pz, nbar)
            ckg = ol.Ckg(self.bg, Pk, L, zmin, zmax, pz, bz, zeff=zeff)
            ckk = ol.Ckk(self.bg, Pk, L)
        return ({'gg': cgg, 'kg': ckg, 'gk': ckg, 'kk': ckk},
                {'gg': ngg, 'kg': np.zeros_like(ckg), 'gk': np.zeros_like(ckg), 'kk': nkk})


def reference_staged():
    from oracle import refload
    return refload.root() is not None


_REF_NKK = None


def cpu_n0(Ls, cfg, inputs, reference=False):
    """N0 on the CPU: the unmodified reference's cmb/Nkk.py (`reference`) or the numpy port oracle/n0.py."""
    from oracle import n0 as on0
    ell, lens, nolens = inputs
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    if reference:
        global _REF_NKK
        if _REF_NKK is None:
            from oracle import refload
            _REF_NKK = refload.load('Nkk')
        return _REF_NKK.Nkk(li, ni, np.asarray(Ls), list(TERMS), cfg[0], cfg[1], iterative=cfg[2], pickleit=False)
    return on0.nkk_root(li, ni, Ls, tuple(TERMS), cfg[0], cfg[1], cfg[2])


def cpu_arm_rate(Ls, cfg, inputs, fc, reference):
    """L/s of the CPU arm on ONE core for the sample `Ls`: N0 and the forecast stage timed separately; the P(k,z) rows of the
    forecast stage are filled by an untimed first pass (the reference builds its camb interpolator once, outside its loops).
    Returns (rate, seconds in N0, seconds in the forecast stage)."""
    k, p, bands = fc
    of = OracleForecast(k, p, reference=reference)
    Lf = Ls.astype(np.float64)
    forecast_stage((of, of), of.Pk, bands[:1], Lf[:1], np.ones(1), np.ones(1))     # warm: every z node of the three grids
    t = time.perf_counter()
    nkk = cpu_n0(Ls, cfg, inputs, reference=reference)
    t1 = time.perf_counter()
    forecast_stage((of, of), of.Pk, bands, Lf, np.ones_like(Lf), nkk)
    t2 = time.perf_counter()
    return len(Ls) / (t2 - t), t1 - t, t2 - t1


def _ref_worker(args):
    L, cfg, reference = args
    return cpu_n0(np.array([L]), cfg, make_inputs(), reference=reference)[0]


def run_reference(a, rank, world):
    """CPU arm on all host cores.  With the reference staged (oracle/_ref/refpy.zip, or /root/reference in the build container) N0 is the
    UNMODIFIED cmb/Nkk.py and the spectra are the UNMODIFIED Cgg.py / Ckg.py / cmb/lensing.py (`kind: reference`); the covariance /
    Fisher / S/N sums (negligible time) are the numpy restatement of the interloper.py script body, which is not importable as a
    function.  Otherwise the numpy port (`kind: port`).  One step = N0 for `nper` L values spread over the grid, one per worker
    process, then the Limber / S/N / Fisher stage on those same L values in the parent (a bounded sample of the b200 arm's step)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    cfg = SWEEP[0]
    reference = reference_staged() and not a.ref_port
    Lgrid = np.arange(a.lmin, a.lmax + 1)
    nper = min(cores, 32)
    sample = np.unique(np.linspace(0, Lgrid.size - 1, nper).astype(int))
    Ls = Lgrid[sample]
    k, p, bands = forecast_inputs()
    of = OracleForecast(k, p, reference=reference)
    ctxmp = mp.get_context('fork')
    t_n0 = t_fc = 0.0

    def one_step(pool):
        t0 = time.perf_counter()
        nkk = np.array(pool.map(_ref_worker, [(int(L), cfg, reference) for L in Ls]))
        t1 = time.perf_counter()
        Lf = Ls.astype(np.float64)
        forecast_stage((of, of), of.Pk, bands, Lf, np.ones_like(Lf), nkk)
        return t1 - t0, time.perf_counter() - t1

    with ctxmp.Pool(len(Ls)) as pool:
        for _ in range(1 if a.warmup > 0 else 0):      # one warm-up pass pages numpy/scipy in and fills the P(k,z) rows
            one_step(pool)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            d0, d1 = one_step(pool)
            t_n0 += d0
            t_fc += d1
        dt = time.perf_counter() - t0
    rate = len(Ls) * a.steps / dt
    kind = 'reference' if reference else 'port'
    what = ('UNMODIFIED reference: cmb/Nkk.py, Cgg.py, Ckg.py, cmb/lensing.py (oracle/refload.py); covariance / Fisher sums: oracle/fisher.py'
            if reference else 'numpy port: oracle/n0.py, oracle/limber.py, oracle/fisher.py')
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': 'L-modes/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': dt / a.steps * 1e3,
        'higher_is_better': True, 'scaling': 'strong' if a.gpus > 1 else 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': WORKLOAD % (cfg[0], cfg[1], a.lmin, a.lmax),
                   'sample': '%d L values per step spread over the grid: N0 one L per worker process, forecast stage in the parent' % len(Ls)},
        'cpu_baseline': {'value': rate, 'unit': 'L-modes/s', 'cores': len(Ls), 'kind': kind,
                         'sample': '%d of %d L per step (%s); N0 %.1f s + forecast %.1f s per step' % (len(Ls), Lgrid.size, what, t_n0 / a.steps, t_fc / a.steps)},
        'e2e': {'value': rate, 'unit': 'L-modes/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    emit(line)


_JSON_FD = None


def _stdout_for_json_only():
    """Everything libraries write to fd 1 (NCCL prints its version banner there, the reference prints its progress) goes to
    stderr; the JSON line alone is written to the original stdout by emit()."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + '\n').encode())


def n0_source_hash():
    h = hashlib.sha256()
    for f in ('n0.cu', 'common.cuh'):
        h.update(open(os.path.join(ROOT, 'lbgcmb_b200', 'csrc', f), 'rb').read())
    return h.hexdigest()[:16]


def load_ncu_profile():
    """The newest committed ncu summary of the dominant kernel (scripts/ncu_raw_summary.py writes it from an `ncu --set full`
    capture of scripts/n0_profile_case.py): executed FP64 flop per unmasked node, pipe utilisations, DRAM bytes per launch, the
    commit and the source hash it was taken at."""
    files = sorted(glob.glob(os.path.join(ROOT, 'profiles', '*_n0_rows5_ncu.json')))
    if not files:
        return None
    prof = json.load(open(files[-1]))
    prof['file'] = os.path.relpath(files[-1], ROOT)
    prof['stale'] = prof.get('source_sha') != n0_source_hash()
    return prof


def main():
    a = parse()
    _stdout_for_json_only()
    rank = int(os.environ.get('RANK', '0'))
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local = int(os.environ.get('LOCAL_RANK', '0'))
    if a.impl == 'reference':
        run_reference(a, rank, world)
        return
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from lbgcmb_b200 import Nkk, _ffi, interloper, pmh, snr as snr_mod, sweep
    from lbgcmb_b200 import dist as ldist
    from lbgcmb_b200.prep_camb import cl_interps_from_tables

    inputs = make_inputs()
    ell, lens, nolens = inputs
    li, ni = cl_interps_from_tables(ell, lens, nolens)
    Ls = np.arange(a.lmin, a.lmax + 1)
    Lf = Ls.astype(np.float64)
    nmodes = np.ones_like(Lf)                        # every integer L is its own bin (prep_Llls.py:13-20 with log10=False)
    cfg = SWEEP[0]
    ctx0, ctx1 = _ffi.context(local), _ffi.context(local, 1)
    fc = forecast_inputs()
    kk, pp, bands = fc
    samples, xsamples = band_samples(bands)
    # P(k) provider = the `Pk_interps` INPUT of Cgg/Ckg/Ckk (pmh.py:18-24; CAMB() in the reference drivers): built once, like
    # the C_l interps above; its device tables are dropped and rebuilt inside every step.
    pk_table = pmh.PkTable(kk, pp, device=local)
    Pk = pmh.get_PkInterps(pk_table)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')   # > 126 MB L2
    stage, timing = {}, {}

    def step(overlap):
        """Host tables / callables in, host results out: ONE call of the public driver.  overlap=False: every library call after the
        previous one on the default context (per-stage device times); overlap=True: N0 enqueued first on the second context."""
        ctx0.kernel_totals(reset=True)
        ctx1.kernel_totals(reset=True)
        pk_table.engine.invalidate_device()          # this step uploads the P(k) table and rebuilds every P(k,z) row
        res = sweep.sharded_forecast(li, ni, Pk, Ls, nmodes, cfg, samples, FSKY, xsample=xsamples, timing=timing, overlap=overlap)
        ms0, n0_ = ctx0.kernel_totals()
        ms1, n1_ = ctx1.kernel_totals()
        stage.update(kernel_ms=ms0 + ms1, launches=n0_ + n1_, n0_ms=Nkk.last_stats['kernel_ms'], main_ms=Nkk.last_stats['main_kernel_ms'],
                     nodes=Nkk.last_stats['nodes'], unmasked=Nkk.last_stats['nodes_unmasked'], path=Nkk.last_stats['path'],
                     n_L_mine=Nkk.last_stats['n_L'], xchg_ms=timing.get('exchange_device_ms') or 0.0)
        if ldist.last_exchange_backend == 'p2p':
            stage['launches'] += 1                   # the peer-memory exchange kernel (csrc/xchg.cu) is ours too
        return res

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step(False)
        res_w = step(True)
    # the fused path against the per-function mirrors (Cgg.Cgg, Ckg.Ckg, lensing.Ckk, interloper.fisher_bias, snr.snr on the gathered N0 curve)
    if rank == 0:
        chk = forecast_stage((interloper, snr_mod), Pk, bands, Lf, nmodes, res_w['nkk'])
        dev = float(np.max(np.abs(rows_of(res_w) / chk - 1)))
        assert dev < 1e-9, 'fused forecast deviates from the per-function mirrors: %.2e' % dev
    peak = ctx0.fp64_peak_tflops() if rank == 0 else 0.0
    sampler = ClockSampler(local)
    sampler.start()
    # ---- timed region 1: K steps, calls one after another: device time per stage ---------------------------------------------------
    barrier()
    acc = {'kernel_ms': 0.0, 'launches': 0, 'n0_ms': 0.0, 'main_ms': 0.0, 'xchg_ms': 0.0}
    seq_s = 0.0
    t_region = time.perf_counter()
    for _ in range(a.steps):
        flush.zero_()                       # L2 flush between timed iterations (tables are a few MB and L2-resident by design)
        barrier()
        t0 = time.perf_counter()
        res_seq = step(False)
        seq_s += time.perf_counter() - t0
        for k_ in acc:
            acc[k_] += stage[k_]
    barrier()
    nodes_step, unmasked_step, path = stage['nodes'], stage['unmasked'], stage['path']
    # ---- timed region 2: K steps as a user runs them (N0 overlapped with the Limber stage): e2e -------------------------------------
    ovl_s = 0.0
    phases = []
    for _ in range(a.steps):
        flush.zero_()
        barrier()
        t0 = time.perf_counter()
        res = step(True)
        torch.cuda.synchronize()
        ovl_s += time.perf_counter() - t0
        phases.append(dict(timing))
    barrier()
    t_region = time.perf_counter() - t_region
    clocks = sampler.summary()
    assert path == 5, 'the bench workload must run the range-walking kernel, got path %s' % path
    for key in ('nkk', 'ckk', 'cgg', 'ckg'):
        assert np.all(np.isfinite(res[key])) and np.array_equal(res[key], res_seq[key]), key
    assert np.all(res['nkk'] > 0) and np.all(np.isfinite(rows_of(res)))
    assert np.max(np.abs(rows_of(res) / rows_of(res_seq) - 1)) < 1e-13, 'overlapped step changed the results'

    # device time of the step = max over ranks of the summed kernel times + the exchange.  The exchange is timed with CUDA events on
    # every rank; a rank that arrives early also waits for its peers inside the collective, so the time of the exchange itself is the
    # MIN over ranks (the last rank to arrive sees no wait).  The skew is host time, which e2e (wall clock) contains.
    t = torch.tensor([acc['kernel_ms'], seq_s, ovl_s, acc['n0_ms'], acc['main_ms'], -acc['xchg_ms']], dtype=torch.float64, device='cuda')
    counts = torch.tensor([float(nodes_step), float(unmasked_step), float(acc['launches'])], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(counts, op=dist.ReduceOp.SUM)
    kern_ms, seq_s, ovl_s, n0_ms, main_ms_sum, xchg_ms = (float(x) for x in t)
    xchg_ms = -xchg_ms
    dev_ms = kern_ms + xchg_ms
    nodes_all, unmasked_all, launches_all = (float(x) for x in counts)
    # the exchange alone on the last step's payload, both routes (device ms, median of 20; collective calls: every rank takes part)
    xchg_how, xchg_cmp = 'none', None
    if world > 1:
        backend = ldist.last_exchange_backend
        xs, xc = np.ones((len(bands), 9)), np.ones((2 + 2 * len(bands), len(range(rank, Ls.size, world))))
        def xtime():
            t_ = []
            for _ in range(20):
                barrier()
                t_.append(ldist.exchange(xs, xc, Ls.size)[2])
            return float(np.median(t_))
        ms_used = xtime()
        xchg_cmp = {'backend': backend, 'device_ms': ms_used, 'p2p_unavailable_because': ldist._p2p['why']}
        if backend == 'p2p':
            keep_ctx, ldist._p2p['ctx'] = ldist._p2p['ctx'], None
            xchg_cmp['nccl_device_ms'] = xtime()
            ldist._p2p['ctx'] = keep_ctx
            xchg_how = ('over NVLink peer memory: ONE kernel per rank stores its block into every peer\'s window and waits for theirs (csrc/xchg.cu), sums '
                        'added in rank order; the NCCL route (all_reduce + all_gather_into_tensor) of the same payload is timed beside it in `exchange`')
        else:
            xchg_how = 'one NCCL all_reduce + one all_gather_into_tensor'
    # per-rank phase breakdown of the last overlapped step (wall ms on that rank)
    mine = {k_: (v * 1e3 if k_.endswith('_s') else v) for k_, v in phases[-1].items()}
    mine.update(rank=rank, kernel_ms_seq=stage['kernel_ms'], n0_kernel_ms=stage['n0_ms'])
    per_rank = [mine]
    if world > 1:
        per_rank = [None] * world
        dist.all_gather_object(per_rank, mine)

    # ---- replica rate (what round 1 reported as its multi-GPU line): rank r runs the full grid for sweep configuration r ----------------
    sweep_rate = None
    if world > 1 and not a.no_replica_rate:
        rcfg = SWEEP[rank % len(SWEEP)]
        barrier()
        rep_ms = 0.0
        for _ in range(2):
            Nkk.n0_terms(li, ni, Ls, TERMS, rcfg[0], rcfg[1], rcfg[2], device=local)
            rep_ms += Nkk.last_stats['kernel_ms']
        tt = torch.tensor([rep_ms], dtype=torch.float64, device='cuda')
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        sweep_rate = Ls.size * world * 2 / (float(tt[0]) * 1e-3)

    units = Ls.size * a.steps                      # strong scaling: the job is ONE grid per step, whatever N
    value = units / (dev_ms * 1e-3)
    e2e = units / ovl_s
    if rank == 0:
        main_ms = main_ms_sum / a.steps             # n0_rows5_kernel alone, max over ranks, per step (CUDA events around the launch)
        share = main_ms * a.steps / dev_ms
        # every rank's kernel sees its own shard: roofline on rank-summed nodes over the slowest rank's kernel time x ranks
        conv = unmasked_all * OPS_PER_NODE / (main_ms * 1e-3) * 1e-12 / world
        prof = load_ncu_profile()
        fpn = prof.get('flop_per_unmasked_node') if prof else None
        ach = unmasked_all * fpn / (main_ms * 1e-3) * 1e-12 / world if fpn else None
        n_band = len(bands)
        fc_ms = (kern_ms - n0_ms) / a.steps
        # bytes counted from the arrays handed to / returned by the C ABI per step on one rank (pinned staging blocks)
        nLr = int(np.ceil(Ls.size / world))
        nz = (1000, 999, 2000)                                   # Cgg, Ckg, Ckk z nodes (interloper.py:148-149, cmb/lensing.py:27)
        h2d = 8 * (nLr                                                               # N0: L (grids / tables only when they change)
                   + 3 * kk.size + 3 * sum(8 * z for z in nz)                        # P(k) table, weights; halo rows of the three grids
                   + nLr + 2 * sum(nz) + 3 * nz[0] + 3 * nz[1] + nz[2] + 2 * 2 * n_band * (nz[0] + nz[1]) + 2 * n_band   # forecast_spectra block
                   + 3 * nLr + 4 * n_band * nLr + nLr + 8 * n_band)                  # forecast_grid block
        d2h = 8 * (7 * nLr + 500 + (4 * n_band + 1) * nLr + 8 * n_band)
        mp_ = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.isfile(os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {}
        line = {
            'metric': METRIC, 'value': value, 'unit': 'L-modes/s', 'n_gpus': world,
            'steps': a.steps, 'warmup': max(a.warmup, 3), 'ms_per_step': dev_ms / a.steps,
            'higher_is_better': True, 'scaling': 'strong' if world > 1 else 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': WORKLOAD % (cfg[0], cfg[1], a.lmin, a.lmax), 'n_L': int(Ls.size), 'call': 'lbgcmb_b200.sweep.sharded_forecast',
                       'phi_fold': True, 'nodes_walked_per_step': nodes_all, 'nodes_unmasked_per_step': unmasked_all,
                       'multi_gpu': ('ONE forecast per step, multipole grid sharded L[r::N]; the exchange inside the timed region -- all-reduce of %d partial '
                                     'sums + all-gather of %d curves -- %s' % (9 * n_band, 2 + 2 * n_band, xchg_how)) if world > 1 else 'single GPU',
                       'cl_tables': 'handed over every step; the library compares them with its resident copy and re-solves / re-uploads only on change',
                       'pk_tables': 'linear P(k) uploaded and every P(k,z) row (3 z grids) rebuilt every step',
                       'l2': 'flushed between timed iterations (256 MB memset); tables are re-read from HBM each step'},
            'rates': {'n0_L_per_s': units / (n0_ms * 1e-3),
                      'limber_cl_per_s': 2 * n_band * 3 * units / a.steps / (fc_ms * 1e-3),
                      'fisher_modes_per_s': n_band * units / a.steps / (fc_ms * 1e-3),
                      'n0_kernel_ms_per_step': n0_ms / a.steps, 'forecast_kernel_ms_per_step': fc_ms, 'exchange_ms_per_step': xchg_ms / a.steps,
                      'sweep_replica_L_per_s': sweep_rate,
                      'sweep_replica_note': 'rank r = full L grid for sweep configuration r (beam/noise), no collective: the round-1 multi-GPU line' if sweep_rate else None},
            'e2e': {'value': e2e, 'unit': 'L-modes/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h, 'ms_per_step': ovl_s / a.steps * 1e3,
                    'how': 'K further steps after the device-timed ones: same host inputs in / host results out through sweep.sharded_forecast, N0 enqueued '
                           "first on the library's second context (lbg_n0_ho02_begin) and joined before the variances; wall clock, barrier on both sides of "
                           'every step, max over ranks; results asserted identical to the sequential steps',
                    'host_forecast_ms_per_step': float(np.mean([(ph_['spectra_s'] + ph_['grid_s']) * 1e3 for ph_ in phases])),
                    'host_forecast_note': 'rank 0, mean over the K overlapped steps: wall time of the spectra + sums part of the step (N0 enqueue, table '
                                          'upload, p(z) / b(z) callables, lbg_forecast_spectra incl. its own kernels, lbg_forecast_grid)',
                    'sequential': {'value': units / seq_s, 'ms_per_step': seq_s / a.steps * 1e3,
                                   'how': 'the K device-timed steps, library calls strictly one after another'}},
            'gpu_launches': int(launches_all),
            'exchange': xchg_cmp,
            'phases_per_rank_ms': per_rank,
            'roofline': {'bound': 'fp64', 'achieved': conv, 'peak': peak, 'unit': 'TFLOP/s', 'frac': conv / peak if peak else None,
                         'definition': 'SURVEY.md 8(d): 145 FP64 op per integrand node x nodes that pass a mask (counted by the kernel: %d per step) / kernel '
                                       'time measured live with CUDA events / measured DFMA peak' % int(unmasked_all),
                         'achieved_executed': ach, 'frac_executed': ach / peak if (ach and peak) else None,
                         'executed_definition': 'FP64 flop the kernel executes per unmasked node (2 DFMA + DMUL + DADD thread instructions of the ncu capture / its '
                                                'unmasked nodes) x unmasked nodes / kernel time',
                         'ncu_profile': {k_: prof.get(k_) for k_ in ('file', 'commit', 'source_sha', 'flop_per_unmasked_node', 'pipes', 'nodes_unmasked', 'kernel')} if prof else None,
                         'source_commit': prof.get('commit') if prof else None,
                         'stale': (prof['stale'] if prof else True), 'source_sha_now': n0_source_hash(),
                         'traffic': prof.get('dram_bytes_per_launch') if prof else None,
                         'traffic_note': 'ncu dram__bytes_read.sum + dram__bytes_write.sum of the profiled launch: the spline tables once; algorithmic bytes per launch = 8 B x n_L output',
                         'peak_source': 'measured DFMA micro-benchmark in this run (lbg_fp64_peak); nominal %.1f; MEASURED_PEAKS.json has no FP64 entry '
                                        '(hbm_gbs=%s is not the bound: arithmetic intensity > 1e4 flop/B)' % (FP64_NOMINAL_TFLOPS, mp_.get('hbm_gbs')),
                         'kernel': 'n0_rows5_kernel', 'kernel_ms': main_ms, 'kernel_share_of_step': share},
            'clocks': clocks, 'wall_s_timed_region': t_region,
        }
        if not a.no_cpu_baseline and world == 1:                   # rank 0 at N = 1 only: the other ranks would idle in the final barrier
            sample = np.array([50, 162, 520, 1677, 3660, 5000][:max(a.cpu_sample, 1)])     # SURVEY.md 8(d) subset
            sample = sample[(sample >= a.lmin) & (sample <= a.lmax)]
            reference = reference_staged()
            r, s_n0, s_fc = cpu_arm_rate(sample, SWEEP[0], inputs, fc, reference)
            line['cpu_baseline'] = {'value': r, 'unit': 'L-modes/s', 'cores': 1, 'kind': 'reference' if reference else 'port',
                                    'sample': 'L=%s of the %d-L grid, 1 process: N0 %.1f s (%s), forecast stage %.1f s (%s; P(k,z) rows warm)' % (
                                        sample.tolist(), Ls.size, s_n0, 'unmodified cmb/Nkk.py' if reference else 'oracle/n0.py', s_fc,
                                        'unmodified Cgg.py / Ckg.py / cmb/lensing.py + oracle/fisher.py' if reference else 'oracle/limber.py + oracle/fisher.py'),
                                    'n0_L_per_s': len(sample) / s_n0, 'forecast_L_per_s': len(sample) / s_fc}
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
