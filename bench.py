#!/usr/bin/env python
"""bench.py -- headline benchmark of the LBGCMB forecast hot path on B200 (see DESIGN.md "Measurement").

One step = the Hu-Okamoto N0 (TT+TE+EE+EB, iterative EB; BASELINE.json configs[1]: thetab=1.4', DeltaT=1 uK',
every integer L in [50, 5000] -> 4951 L-modes) followed by the Limber C_kk/C_kg/C_gg + SNR/Fisher stage on the
same multipoles when that stage is built (--no-limber to leave it out).  metric = L-modes per second.

  value      whole-job L-modes/s, device time of the kernels (CUDA events on the library's stream), tables resident
  e2e        the same metric through the reference-facing Python call (Nkk.Nkk(...)) with HOST inputs: spline build,
             table upload, H2D, kernels, D2H all inside the timed region
  roofline   dominant kernel (n0_kernel): FP64-pipe bound, achieved = executed integrand nodes x 145 op / time
  cpu_baseline  the numpy port of the reference (oracle/n0.py) on this box's host cores, bounded sample

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
FP64_NOMINAL_TFLOPS = 37.2  # 148 SM x 64 DFMA/clk x 2 x 1.965 GHz

# (thetab [arcmin], DeltaT [uK-arcmin], iterative): rank r of a multi-GPU run takes SWEEP[r % len]
SWEEP = [(1.4, 1.0, True), (1.4, 7.0, False), (1.5, 12.0, False), (7.0, 30.0, False),
         (1.0, 1.0, True), (2.0, 2.0, True), (3.0, 5.0, False), (5.0, 20.0, False)]
TERMS = ['TT', 'TE', 'EE', 'EB']


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=5)
    ap.add_argument('--warmup', type=int, default=3)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--lmin', type=int, default=50)
    ap.add_argument('--lmax', type=int, default=5000)
    ap.add_argument('--no-fold', action='store_true', help='evaluate all 201 phi nodes instead of 101 mirror pairs')
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


def cpu_port_rate(Ls, cfg, inputs):
    """L/s of the numpy port of cmb/Nkk.py (oracle/n0.py) on one core."""
    from oracle import n0 as on0
    ell, lens, nolens = inputs
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    t = time.perf_counter()
    on0.nkk_root(li, ni, Ls, tuple(TERMS), cfg[0], cfg[1], cfg[2])
    return len(Ls) / (time.perf_counter() - t)


def _ref_worker(args):
    L, cfg = args
    from oracle import n0 as on0
    ell, lens, nolens = make_inputs()
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    return on0.nkk_root(li, ni, np.array([L]), tuple(TERMS), cfg[0], cfg[1], cfg[2])[0]


def run_reference(a, rank, world):
    """CPU arm: the numpy port of the reference (oracle/n0.py; /root/reference is absent on the GPU box) on all
    host cores, one L per worker per step (bounded sample of the same L grid and configuration)."""
    if rank != 0:
        return
    import multiprocessing as mp
    cores = os.cpu_count() or 1
    cfg = SWEEP[0]
    Lgrid = np.arange(a.lmin, a.lmax + 1)
    nper = min(cores, 32)
    sample = [int(x) for x in np.linspace(0, Lgrid.size - 1, nper).astype(int)]
    ctxmp = mp.get_context('fork')
    with ctxmp.Pool(nper) as pool:
        jobs = [(int(Lgrid[i]), cfg) for i in sample]
        for _ in range(max(a.warmup, 0) and 1):      # one warm-up pass is enough to page numpy/scipy in
            pool.map(_ref_worker, jobs)
        t0 = time.perf_counter()
        for _ in range(a.steps):
            pool.map(_ref_worker, jobs)
        dt = time.perf_counter() - t0
    rate = nper * a.steps / dt
    line = {
        'impl': 'reference', 'metric': 'N0+Limber L-modes/sec (FP64)', 'value': rate, 'unit': 'L-modes/s',
        'n_gpus': a.gpus, 'steps': a.steps, 'warmup': a.warmup, 'ms_per_step': dt / a.steps * 1e3,
        'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
        'config': {'workload': 'N0 Nkk TT/TE/EE/EB iterative, thetab=1.4, DeltaT=1.0, L=%d..%d' % (a.lmin, a.lmax),
                   'sample': '%d L values per step spread over the grid, one per worker' % nper},
        'cpu_baseline': {'value': rate, 'unit': 'L-modes/s', 'cores': nper, 'kind': 'port',
                         'sample': '%d of %d L per step (numpy port of cmb/Nkk.py, oracle/n0.py)' % (nper, Lgrid.size)},
        'e2e': {'value': rate, 'unit': 'L-modes/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line), flush=True)


def main():
    a = parse()
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
    from lbgcmb_b200 import Nkk, _ffi
    from lbgcmb_b200.prep_camb import cl_interps_from_tables

    inputs = make_inputs()
    ell, lens, nolens = inputs
    li, ni = cl_interps_from_tables(ell, lens, nolens)
    Ls = np.arange(a.lmin, a.lmax + 1)
    cfg = SWEEP[rank % len(SWEEP)]
    fold = not a.no_fold
    ctx = _ffi.context(local)
    flush = torch.empty(256 << 20, dtype=torch.uint8, device='cuda')   # > 126 MB L2

    def step():
        return Nkk.n0_terms(li, ni, Ls, TERMS, cfg[0], cfg[1], cfg[2], fold_phi=fold, device=local)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(max(a.warmup, 3)):
        step()
    peak = ctx.fp64_peak_tflops() if rank == 0 else 0.0
    sampler = ClockSampler(local)
    sampler.start()
    barrier()
    kernel_ms, e2e_s, launches, nodes = 0.0, 0.0, 0, 0
    t_region = time.perf_counter()
    for _ in range(a.steps):
        flush.zero_()                       # L2 flush between timed iterations (tables are 1.3 MB and L2-resident by design)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        out, _terms = step()                # host tables in, host N0 out: e2e
        e2e_s += time.perf_counter() - t0
        kernel_ms += Nkk.last_stats['kernel_ms']
        launches += Nkk.last_stats['launches']
        nodes = Nkk.last_stats['nodes']
    barrier()
    t_region = time.perf_counter() - t_region
    clocks = sampler.summary()
    assert np.all(np.isfinite(out)) and np.all(out > 0)

    t = torch.tensor([kernel_ms, e2e_s], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    kernel_ms, e2e_s = float(t[0]), float(t[1])
    units = Ls.size * world * a.steps
    value = units / (kernel_ms * 1e-3)
    e2e = units / e2e_s
    if rank == 0:
        ach = nodes * OPS_PER_NODE * a.steps / (kernel_ms * 1e-3) * 1e-12
        h2d = 8 * (2 * 4 * ell.size + Ls.size + 2 * 9999 + 3 * 201) + 4 * 9999
        line = {
            'metric': 'N0+Limber L-modes/sec (FP64)', 'value': value, 'unit': 'L-modes/s', 'n_gpus': world,
            'steps': a.steps, 'warmup': max(a.warmup, 3), 'ms_per_step': kernel_ms / a.steps,
            'higher_is_better': True, 'scaling': 'weak', 'vs_baseline': None, 'dtype': 'f64', 'data': 'synthetic',
            'config': {'workload': 'N0 Nkk TT/TE/EE/EB iterative, thetab=1.4, DeltaT=1.0, L=%d..%d (%d L-modes x 9999 l1 x 201 phi)'
                                   % (a.lmin, a.lmax, Ls.size),
                       'phi_fold': fold, 'nodes_executed_per_step': nodes, 'multi_gpu': 'rank r = sweep config r, full L grid each',
                       'l2': 'flushed between timed iterations (256 MB memset); tables (1.3 MB) are re-read from HBM each step'},
            'e2e': {'value': e2e, 'unit': 'L-modes/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': 8 * 6 * Ls.size,
                    'ms_per_step': e2e_s / a.steps * 1e3},
            'gpu_launches': launches,
            'roofline': {'bound': 'fp64', 'achieved': ach, 'peak': peak, 'unit': 'TFLOP/s', 'frac': ach / peak if peak else None,
                         'traffic': None, 'peak_source': 'measured DFMA micro-benchmark (lbg_fp64_peak); nominal %.1f' % FP64_NOMINAL_TFLOPS,
                         'ops_per_node': OPS_PER_NODE, 'kernel': 'n0_kernel'},
            'clocks': clocks, 'wall_s_timed_region': t_region,
        }
        if not a.no_cpu_baseline:
            sample = np.linspace(a.lmin, a.lmax, a.cpu_sample).astype(int)
            r = cpu_port_rate(sample, SWEEP[0], inputs)
            line['cpu_baseline'] = {'value': r, 'unit': 'L-modes/s', 'cores': 1, 'kind': 'port',
                                    'sample': 'L=%s of the %d-L grid, numpy port of cmb/Nkk.py (oracle/n0.py), 1 process'
                                              % (sample.tolist(), Ls.size)}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
