"""profiles/<tag>_n0_rows5_ncu.json from an `ncu --set full` capture of scripts/n0_profile_case.py: what bench.py's roofline reads
instead of hard-coded constants (executed FP64 flop per unmasked node, pipe utilisations, DRAM bytes per launch), tied to the
commit and to the hash of the kernel source it was taken at.

  ncu -i gpurun_out/<tag>_n0_rows5.ncu-rep --page raw --csv > raw.csv
  python scripts/ncu_n0_json.py raw.csv gpurun_out/<tag>_ncu_full.log <tag>        (the log holds the case's JSON line)"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main(raw, caselog, tag):
    import bench
    rows = list(csv.reader(open(raw)))
    hdr, vals = rows[0], rows[2]
    m = dict(zip(hdr, vals))

    def f(name):
        return float(m[name].replace(',', ''))
    case = None
    for line in open(caselog):
        if line.startswith('{') and 'n0_profile_case' in line:
            case = json.loads(line)
    if case is None:
        raise SystemExit('no n0_profile_case line in %s' % caselog)
    cycles = f('sm__cycles_elapsed.avg')
    per_clk = {k: f('smsp__sass_thread_inst_executed_op_%s_pred_on.sum.per_cycle_elapsed' % k) for k in ('dadd', 'dmul', 'dfma')}
    flop = (per_clk['dadd'] + per_clk['dmul'] + 2. * per_clk['dfma']) * cycles
    out = {
        'tag': tag, 'kernel': m['Kernel Name'],
        'commit': subprocess.run(['git', '-C', ROOT, 'rev-parse', '--short', 'HEAD'], capture_output=True, text=True).stdout.strip(),
        'source_sha': bench.n0_source_hash(),
        'case': case,
        'nodes_unmasked': case['nodes_unmasked'], 'nodes_walked': case['nodes_walked'],
        'gpu_time_ms': f('gpu__time_duration.sum'),
        'fp64_flop_executed': flop,
        'flop_per_unmasked_node': flop / case['nodes_unmasked'],
        'thread_inst_per_unmasked_node': {k: v * cycles / case['nodes_unmasked'] for k, v in per_clk.items()},
        'dram_bytes_per_launch': f('dram__bytes_read.sum') * {'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'Gbyte': 1e9}[rows[1][hdr.index('dram__bytes_read.sum')]]
        + f('dram__bytes_write.sum') * {'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1.0, 'Gbyte': 1e9}[rows[1][hdr.index('dram__bytes_write.sum')]],
        'pipes': {'fp64_pipe_active': f('sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active') / 100.,
                  'lsu_data_pipe_wavefronts': f('l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed') / 100.,
                  'issue_active': f('smsp__issue_active.avg.pct_of_peak_sustained_active') / 100.},
        'registers_per_thread': f('launch__registers_per_thread'),
        'lsu_wavefronts_per_unmasked_warp_node': f('l1tex__data_pipe_lsu_wavefronts.sum') / (case['nodes_unmasked'] / 32.) if 'l1tex__data_pipe_lsu_wavefronts.sum' in m else None,
    }
    path = os.path.join(ROOT, 'profiles', '%s_n0_rows5_ncu.json' % tag)
    json.dump(out, open(path, 'w'), indent=1)
    print(path, json.dumps(out)[:400])


if __name__ == '__main__':
    main(*sys.argv[1:4])
