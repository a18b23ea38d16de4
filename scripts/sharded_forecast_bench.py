"""One forecast (N0 + Limber + S/N + Fisher for one experiment and one dropout sample) with the 4951-multipole grid sharded over N GPUs
and a single NCCL all-reduce of the partial Fisher sums (sweep.sharded_forecast): strong scaling of the path north_star describes.

  python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 scripts/sharded_forecast_bench.py

Prints one JSON line on rank 0: wall time per forecast (max over ranks, barrier on both sides), the results, and their deviation from
the values a single rank obtains for the whole grid (computed by rank 0 at the end, outside the timed region)."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def main():
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (('RANK', '0'), ('WORLD_SIZE', '1'), ('LOCAL_RANK', '0')))
    import bench
    bench._stdout_for_json_only()                       # before NCCL initialises: its version banner goes to fd 1
    import torch
    import torch.distributed as dist
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    from lbgcmb_b200 import pmh, sweep
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    li, ni = cl_interps_from_tables(*bench.make_inputs())
    kk, pp, bands = bench.forecast_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp, device=local))
    Ls = np.arange(50, 5001)
    nm = np.ones(Ls.size)
    b = bands[0]
    smp = {'pz': b['pz'], 'bz': b['bz'], 'nbar': b['nbar'], 'fid_b1': b['fid_b1'], 'LllMax': b['LllMax']}
    xsmp = {'pz': b['pzz'], 'bz': b['bz'], 'nbar': b['nbar_wint']}

    timing = {}

    def run():
        return sweep.sharded_forecast(li, ni, Pk, Ls, nm, (1.4, 1.0, True), smp, bench.FSKY, xsample=xsmp, timing=timing)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()

    for _ in range(3):
        run()
    reps = 10
    barrier()
    t0 = time.perf_counter()
    for _ in range(reps):
        res = run()
    barrier()
    dt = torch.tensor([(time.perf_counter() - t0) / reps], dtype=torch.float64, device='cuda')
    if world > 1:
        dist.all_reduce(dt, op=dist.ReduceOp.MAX)
    phases = dict(timing)
    if rank == 0:
        dev = None
        if world > 1:                                    # the whole grid on this rank alone, outside any process group
            from lbgcmb_b200 import dist as ldist
            real = ldist._dist
            ldist._dist = lambda: None
            one = run()
            ldist._dist = real
            dev = {'nkk': float(np.max(np.abs(res['nkk'] / one['nkk'] - 1))), 'ckg': float(np.max(np.abs(res['ckg'] / one['ckg'] - 1))),
                   'snr': float(abs(res['snr'] / one['snr'] - 1)), 'Fisher': float(np.max(np.abs(res['Fisher'] / one['Fisher'] - 1))),
                   'dtheta': float(np.max(np.abs(res['dtheta'] / one['dtheta'] - 1)))}
        bench.emit({'workload': 'one forecast, 4951 multipoles sharded over the ranks: N0 (TT,TE,EE,EB iterative) + C_kk/C_gg/C_kg (clean + contaminated) '
                                '+ S/N + (sigma8,b1) Fisher / bias; one NCCL all-reduce of 7 partial sums + curve gathers',
                    'n_gpus': world, 'ms_per_forecast': float(dt[0]) * 1e3, 'L_modes_per_s': Ls.size / float(dt[0]), 'scaling': 'strong',
                    'snr': res['snr'], 'Fisher': res['Fisher'].tolist(), 'dtheta': res['dtheta'].tolist(),
                    'max_rel_dev_vs_one_rank': dev, 'rank0_phase_ms': {k: (v * 1e3 if k.endswith('_s') else v) for k, v in phases.items()}})
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == '__main__':
    main()
