#!/usr/bin/env python
"""bench.py -- headline benchmark of the LBGCMB forecast hot path on B200.

One step = the Hu-Okamoto N0 (TT+TE+EE+EB, iterative EB; BASELINE.json configs[1]: thetab=1.4', DeltaT=1 uK',
every integer L in [50, 5000] -> 4951 L-modes) followed by the Limber C_kk/C_kg/C_gg + covariance + S/N + (sigma8, b1)
Fisher / interloper-bias stage for the u, g, r, i dropout samples on the same multipoles (configs[0], [2]; --no-limber
leaves it out).  metric = L-modes per second.

  value      whole-job L-modes/s over the device time of every kernel of the step (CUDA events on the library's stream)
  e2e        the same metric through the reference-facing Python calls (Nkk.n0_terms, interloper.get_allCls, snr.snr, ...)
             with HOST inputs: spline/table builds, H2D, kernels, D2H all inside the timed region
  roofline   dominant kernel (n0_rows5_kernel): FP64-pipe bound; achieved = executed FP64 flop (ncu count per node) / time
  cpu_baseline  the numpy port of the reference (oracle/) on this box's host cores, bounded sample

N>1 (torchrun, one rank per GPU): weak scaling -- rank r runs the full L grid for sweep configuration r of
BASELINE.json configs[4] (beam/noise sweep); no data-path collective; max-over-ranks timing.
`--impl reference` times the CPU port with all host cores (rank 0 only).
"""
import argparse
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
N0_DRAM_BYTES_PER_LAUNCH = 2212352   # ncu --set full, profiles/r1d_n0_rows5_ncu_full_summary.txt: dram read+write of the 496-L profile launch (the tables; partial sums stay in L2)
# FP64 flop the rows5 kernel EXECUTES per folded quadrature node (2 DFMA + DMUL + DADD thread instructions of the ncu capture
# profiles/r1d_n0_rows5_ncu_full_summary.txt: 7357.9 flop/clk x 4 778 091 clk / 5.011e8 nodes); masked nodes are not visited
EXEC_FLOP_PER_NODE = 70.2
NCU_PIPES = {'fp64_pipe_active': 0.483, 'lsu_data_pipe_wavefronts': 0.769, 'issue_active': 0.498,
             'source': 'profiles/r1d_n0_rows5_ncu_full_summary.txt (ncu --set full, every 10th L of the bench grid)'}
FP64_NOMINAL_TFLOPS = 37.2  # 148 SM x 64 DFMA/clk x 2 x 1.965 GHz

# (thetab [arcmin], DeltaT [uK-arcmin], iterative): rank r of a multi-GPU run takes SWEEP[r % len]
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
    ap.add_argument('--no-fold', action='store_true', help='evaluate all 201 phi nodes instead of 101 mirror pairs')
    ap.add_argument('--no-limber', action='store_true', help='N0 only (kernel experiments)')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--cpu-sample', type=int, default=4, help='number of L values timed on the CPU port')
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


def forecast_stage(mods, Pk, bands, Ls, nmodes, nkk):
    """Limber C_kk / C_kg / C_gg (full z integrals, zmin = 0.01, zmax = 10 as interloper.py:148-149), Gaussian covariance,
    cumulative S/N of C_kg and the (sigma8, b1) Fisher matrix + interloper bias for every dropout band: the body of
    interloper.py:168-260 and snr.py:118-124 through the same-named functions.  `mods` = (interloper, snr) modules of the
    implementation under test (lbgcmb_b200 mirrors, or the oracle for the CPU legs).  Returns one row per band.
    `nkk` is the N0 curve or a Future of it (Nkk.n0_terms_async), resolved after the spectra are built."""
    interloper, snr = mods
    pending = np.zeros(len(Ls))                  # N_kk enters only through Nls['kk']: filled in once the curve is there
    spectra = []
    for b in bands:
        spectra.append((interloper.get_allCls(Pk, Ls, b['nbar'], FSKY, 0.01, 10.0, b['pz'], b['bz'], zeff=False, nkk=pending),
                        interloper.get_allCls(Pk, Ls, b['nbar_wint'], FSKY, 0.01, 10.0, b['pzz'], b['bz'], zeff=False, nkk=pending)))
    if hasattr(nkk, 'result'):                   # a Future from Nkk.n0_terms_async: the spectra above were built while it ran
        nkk = nkk.result()[0]
    rows = []
    for b, ((Cls, Nls), (xCls, xNls)) in zip(bands, spectra):
        Nls['kk'] = xNls['kk'] = nkk
        vkg = ((Cls['kk'] + Nls['kk']) * (Cls['gg'] + Nls['gg']) + Cls['kg'] ** 2.) / ((2. * Ls + 1.) * FSKY)   # Ckg.py:83-87
        sn = snr.snr(Ls, Cls['kg'], vkg, nmodes, lmax=b['LllMax']) / np.sqrt(FSKY)                                # snr.py:122-123
        fb = interloper.fisher_bias(Ls, nmodes, Cls, Nls, xCls, xNls, FSKY, 0.819, b['fid_b1'], b['LllMax'])
        rows.append(np.concatenate([[sn], fb['Fisher'].ravel(), fb['dtheta']]))
    return np.array(rows)


class OracleForecast(object):
    """CPU legs only: the numpy restatements (oracle/limber.py, oracle/halofit.py, oracle/fisher.py) behind the call surface
    forecast_stage() uses.  P(k,z) rows are cached per redshift, the way the reference's camb interpolator is built once."""

    def __init__(self, k, p):
        from oracle import limber as ol, fisher as ofi, halofit as oh, modi_lensing as om
        from lbgcmb_b200.synthetic import PARAMS
        self.ol, self.ofi = ol, ofi
        self.bg = ol.Background(PARAMS)
        mc = om.ModiCosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=k, plin=p)
        self.Pk = oh.ModiPk(k, p, lambda z: float(mc.Dgrow(z=z)), mc.cosmo, PARAMS['sig_8']).interps()
        self.snr = ofi.snr
        self.fisher_bias = ofi.fisher_bias

    def get_allCls(self, Pk, Llls, nbar, fsky, zmin, zmax, pz, bz, zeff=False, nkk=None):
        ol, L = self.ol, np.asarray(Llls, dtype=np.float64)
        cgg = ol.Cgg(self.bg, Pk, L, zmin, zmax, pz, bz, zeff=zeff)
        ngg = ol.Ngg(L, zmin, zmax, pz, nbar)
        ckg = ol.Ckg(self.bg, Pk, L, zmin, zmax, pz, bz, zeff=zeff)
        ckk = ol.Ckk(self.bg, Pk, L)
        return ({'gg': cgg, 'kg': ckg, 'gk': ckg, 'kk': ckk},
                {'gg': ngg, 'kg': np.zeros_like(ckg), 'gk': np.zeros_like(ckg), 'kk': nkk})


def cpu_n0(Ls, cfg, inputs):
    from oracle import n0 as on0
    ell, lens, nolens = inputs
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    return on0.nkk_root(li, ni, Ls, tuple(TERMS), cfg[0], cfg[1], cfg[2])


def cpu_port_rate(Ls, cfg, inputs, fc):
    """L/s of the numpy port (oracle/n0.py + oracle/limber.py + oracle/fisher.py) on one core: N0 and the forecast stage on
    the sample `Ls`.  Returns (rate, seconds in N0, seconds in the forecast stage)."""
    k, p, bands = fc
    of = OracleForecast(k, p)
    Lf = Ls.astype(np.float64)
    t = time.perf_counter()
    nkk = cpu_n0(Ls, cfg, inputs)
    t1 = time.perf_counter()
    forecast_stage((of, of), of.Pk, bands, Lf, np.ones_like(Lf), nkk)
    t2 = time.perf_counter()
    return len(Ls) / (t2 - t), t1 - t, t2 - t1


def _ref_worker(args):
    L, cfg = args
    return cpu_n0(np.array([L]), cfg, make_inputs())[0]


def run_reference(a, rank, world):
    """CPU arm: the numpy port of the reference (oracle/; /root/reference is absent on the GPU box) on all host cores.
    One step = N0 for `nper` L values spread over the grid, one per worker process, then the Limber / S/N / Fisher stage on
    those same L values in the parent (bounded sample of the b200 arm's workload and configuration)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    cfg = SWEEP[0]
    Lgrid = np.arange(a.lmin, a.lmax + 1)
    nper = min(cores, 32)
    sample = np.unique(np.linspace(0, Lgrid.size - 1, nper).astype(int))
    Ls = Lgrid[sample]
    k, p, bands = forecast_inputs()
    of = OracleForecast(k, p)
    ctxmp = mp.get_context('fork')
    t_n0 = t_fc = 0.0

    def one_step(pool):
        t0 = time.perf_counter()
        nkk = np.array(pool.map(_ref_worker, [(int(L), cfg) for L in Ls]))
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
    line = {
        'impl': 'reference', 'metric': METRIC, 'value': rate, 'unit': 'L-modes/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': dt / a.steps * 1e3,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': WORKLOAD % (cfg[0], cfg[1], a.lmin, a.lmax),
                   'sample': '%d L values per step spread over the grid: N0 one L per worker process, forecast stage in the parent' % len(Ls)},
        'cpu_baseline': {'value': rate, 'unit': 'L-modes/s', 'cores': len(Ls), 'kind': 'port',
                         'sample': '%d of %d L per step (numpy port: oracle/n0.py, oracle/limber.py, oracle/fisher.py); '
                                   'N0 %.1f s + forecast %.1f s per step' % (len(Ls), Lgrid.size, t_n0 / a.steps, t_fc / a.steps)},
        'e2e': {'value': rate, 'unit': 'L-modes/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    emit(line)


_JSON_FD = None


def _stdout_for_json_only():
    """Everything libraries write to fd 1 (NCCL prints its version banner there) goes to stderr; the JSON line alone is
    written to the original stdout by emit()."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    os.write(_JSON_FD if _JSON_FD is not None else 1, (json.dumps(line) + '\n').encode())


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
    from lbgcmb_b200 import Nkk, _ffi, interloper, pmh, snr as snr_mod
    from lbgcmb_b200.prep_camb import cl_interps_from_tables

    inputs = make_inputs()
    ell, lens, nolens = inputs
    li, ni = cl_interps_from_tables(ell, lens, nolens)
    Ls = np.arange(a.lmin, a.lmax + 1)
    Lf = Ls.astype(np.float64)
    nmodes = np.ones_like(Lf)                        # every integer L is its own bin (prep_Llls.py:13-20 with log10=False)
    cfg = SWEEP[rank % len(SWEEP)]
    fold = not a.no_fold
    ctx = _ffi.context(local)
    fc = forecast_inputs()
    kk, pp, bands = fc
    # P(k) provider = the `Pk_interps` INPUT of Cgg/Ckg/Ckk (pmh.py:18-24; CAMB() in the reference drivers): built once, like
    # the C_l interps above; its device tables are dropped and rebuilt inside every step.
    pk_table = pmh.PkTable(kk, pp, device=local)
    Pk = pmh.get_PkInterps(pk_table)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')   # > 126 MB L2
    stage = {}

    def step():
        """Host tables in, host results out (the calls a user of the reference makes): returns (N0[L], forecast rows)."""
        ctx.kernel_totals(reset=True)
        t0 = time.perf_counter()
        nkk, _terms = Nkk.n0_terms(li, ni, Ls, TERMS, cfg[0], cfg[1], cfg[2], fold_phi=fold, device=local)
        stage['n0_ms'], stage['n0_launches'] = ctx.kernel_totals()
        stage['nodes'] = Nkk.last_stats['nodes']
        t1 = time.perf_counter()
        rows = None
        if not a.no_limber:
            pk_table.engine.invalidate_device()                            # this step uploads the P(k) table and rebuilds every P(k,z) row
            rows = forecast_stage((interloper, snr_mod), Pk, bands, Lf, nmodes, nkk)
        stage['all_ms'], stage['launches'] = ctx.kernel_totals()
        stage['host_n0_s'], stage['host_fc_s'] = t1 - t0, time.perf_counter() - t1
        return nkk, rows

    def step_overlapped():
        """Same inputs, same results, same calls, with the N0 curve started first on the library's second context
        (Nkk.n0_terms_async) so the Limber spectra are built while it runs; joined before the variances are formed."""
        fut = Nkk.n0_terms_async(li, ni, Ls, TERMS, cfg[0], cfg[1], cfg[2], fold_phi=fold, device=local)
        if a.no_limber:
            return fut.result()[0], None
        pk_table.engine.invalidate_device()
        rows = forecast_stage((interloper, snr_mod), Pk, bands, Lf, nmodes, fut)
        return fut.result()[0], rows

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step()
        step_overlapped()
    peak = ctx.fp64_peak_tflops() if rank == 0 else 0.0
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    acc = {'n0_ms': 0.0, 'all_ms': 0.0, 'launches': 0, 'n0_launches': 0, 'host_n0_s': 0.0, 'host_fc_s': 0.0}
    e2e_s = 0.0
    t_region = time.perf_counter()
    for _ in range(a.steps):
        flush.zero_()                       # L2 flush between timed iterations (tables are a few MB and L2-resident by design)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out, rows = step()                  # e2e: spline/table builds, H2D, kernels, D2H all inside
        e2e_s += time.perf_counter() - t0
        for k_ in acc:
            acc[k_] += stage[k_]
    barrier()
    t_region = time.perf_counter() - t_region
    main_ms = Nkk.last_stats['main_kernel_ms']   # n0_rows5_kernel alone, last sequential step (CUDA events around the launch)
    nodes_step = stage['nodes']
    # second timed region, K more steps: the same step with N0 overlapped (e2e only; kernel timings above stay un-overlapped)
    ovl_s = 0.0
    for _ in range(a.steps):
        flush.zero_()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out2, rows2 = step_overlapped()
        ovl_s += time.perf_counter() - t0
    barrier()
    clocks = sampler.summary()
    assert np.all(np.isfinite(out)) and np.all(out > 0)
    assert rows is None or np.all(np.isfinite(rows))
    assert np.array_equal(out, out2) and (rows is None or np.array_equal(rows, rows2)), 'overlapped step changed the results'

    t = torch.tensor([acc['all_ms'], e2e_s, acc['n0_ms'], ovl_s], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    all_ms, e2e_s, n0_ms, ovl_s = float(t[0]), float(t[1]), float(t[2]), float(t[3])
    units = Ls.size * world * a.steps
    value = units / (all_ms * 1e-3)
    e2e_seq = units / e2e_s
    e2e = units / ovl_s
    if rank == 0:
        nodes = nodes_step
        ach = nodes * EXEC_FLOP_PER_NODE / (main_ms * 1e-3) * 1e-12       # executed FP64 flop/s (hardware count from ncu x nodes / live time)
        conv = nodes * OPS_PER_NODE / (main_ms * 1e-3) * 1e-12           # SURVEY.md 8(d) frozen convention: 145 op per folded node
        n_band = len(bands)
        fc_ms = (acc['all_ms'] - acc['n0_ms']) / a.steps
        # bytes counted from the arrays handed to / returned by the C ABI per step
        h2d_n0 = 8 * (2 * 4 * ell.size + Ls.size + 2 * 9999 + 3 * 201 + 2 * 201 * 8) + 4 * 9999
        nz = (1000, 999, 2000)                                   # Cgg, Ckg, Ckk z nodes (interloper.py:148-149, cmb/lensing.py:27)
        h2d_fc = 0 if a.no_limber else 8 * 3 * kk.size + 2 * n_band * sum(8 * 8 * z + 8 * (Ls.size + 2 * z) + 4 * z for z in nz) \
            + n_band * 8 * Ls.size * (4 + 4 + 9 + 6 + 3 + 1)
        d2h_fc = 0 if a.no_limber else 2 * n_band * 3 * 8 * Ls.size + n_band * 8 * (16 * Ls.size + 1 + 4 + 2)
        mp = json.load(open(os.path.join(ROOT, 'MEASURED_PEAKS.json'))) if os.path.isfile(os.path.join(ROOT, 'MEASURED_PEAKS.json')) else {}
        line = {
            'metric': METRIC, 'value': value, 'unit': 'L-modes/s', 'n_gpus': world,
            'steps': a.steps, 'warmup': max(a.warmup, 3), 'ms_per_step': all_ms / a.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': WORKLOAD % (cfg[0], cfg[1], a.lmin, a.lmax), 'n_L': int(Ls.size), 'limber_stage': not a.no_limber,
                       'phi_fold': fold, 'nodes_executed_per_step': nodes, 'multi_gpu': 'rank r = sweep config r (beam/noise), full L grid each; no data-path collective',
                       'l2': 'flushed between timed iterations (256 MB memset); tables are re-read from HBM each step'},
            'rates': {'n0_L_per_s': units / (n0_ms * 1e-3),
                      'limber_cl_per_s': None if a.no_limber else 2 * n_band * 3 * Ls.size * world / (fc_ms * 1e-3),
                      'fisher_modes_per_s': None if a.no_limber else n_band * Ls.size * world / (fc_ms * 1e-3),
                      'n0_kernel_ms_per_step': acc['n0_ms'] / a.steps, 'forecast_kernel_ms_per_step': fc_ms,
                      'host_n0_ms_per_step': acc['host_n0_s'] / a.steps * 1e3, 'host_forecast_ms_per_step': acc['host_fc_s'] / a.steps * 1e3},
            'e2e': {'value': e2e, 'unit': 'L-modes/s', 'h2d_bytes_per_step': h2d_n0 + h2d_fc,
                    'd2h_bytes_per_step': 8 * 6 * Ls.size + d2h_fc, 'ms_per_step': ovl_s / a.steps * 1e3,
                    'how': 'K further steps after the kernel-timed ones: same host inputs in / host results out, N0 started first on the '
                           "library's second context (Nkk.n0_terms_async) and joined before the variances; results asserted bit-identical",
                    'sequential': {'value': e2e_seq, 'ms_per_step': e2e_s / a.steps * 1e3,
                                   'how': 'the K kernel-timed steps, calls one after another (Nkk, then the forecast stage)'}},
            'gpu_launches': acc['launches'],
            'roofline': {'bound': 'fp64', 'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak if peak else None,
                         'traffic': N0_DRAM_BYTES_PER_LAUNCH, 'traffic_source': 'ncu --set full of scripts/n0_profile_case.py (every 10th L): the spline tables once; '
                                                                                'algorithmic bytes per launch = 8 B x n_L output',
                         'achieved_definition': 'FP64 flop the kernel executes (%.1f per folded node, DFMA = 2: ncu instruction counts) x nodes / kernel time '
                                                'measured live with CUDA events; fully masked nodes (22%% of the grid) are skipped, DMUL/DADD fill a pipe slot '
                                                'with 1 flop, so the pipe is busier than frac says: see pipes' % EXEC_FLOP_PER_NODE,
                         'pipes_ncu': NCU_PIPES,
                         'convention_145op': {'achieved': conv, 'frac': conv / peak if peak else None,
                                              'note': 'SURVEY.md 8(d) op count of the reference integrand (145 per node) x folded nodes / time: algorithmic, not executed'},
                         'peak_source': 'measured DFMA micro-benchmark in this run (lbg_fp64_peak); nominal %.1f; MEASURED_PEAKS.json has no FP64 entry '
                                        '(hbm_gbs=%s is not the bound: arithmetic intensity > 1e4 flop/B)' % (FP64_NOMINAL_TFLOPS, mp.get('hbm_gbs')),
                         'kernel': 'n0_rows5_kernel', 'kernel_ms': main_ms, 'kernel_share_of_step': main_ms * a.steps / all_ms},
            'clocks': clocks, 'wall_s_timed_region': t_region,
        }
        if not a.no_cpu_baseline:
            sample = np.array([50, 162, 520, 1677, 3660, 5000][:max(a.cpu_sample, 1)])     # SURVEY.md 8(d) subset
            sample = sample[(sample >= a.lmin) & (sample <= a.lmax)]
            r, s_n0, s_fc = cpu_port_rate(sample, SWEEP[0], inputs, fc)
            line['cpu_baseline'] = {'value': r, 'unit': 'L-modes/s', 'cores': 1, 'kind': 'port',
                                    'sample': 'L=%s of the %d-L grid, numpy port (oracle/n0.py %.1f s; oracle/limber.py + fisher.py %.1f s '
                                              'incl. first-use HALOFIT rows), 1 process' % (sample.tolist(), Ls.size, s_n0, s_fc)}
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
