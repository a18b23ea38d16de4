"""Summarise an `ncu -i X.ncu-rep --page raw --csv` dump: python scripts/ncu_raw_summary.py raw.csv [substr ...]"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
hdr, units = rows[0], rows[1]
want = sys.argv[2:] or ['Kernel Name', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
                        'launch__occupancy_limit', 'launch__waves', 'sm__warps_active.avg.pct_of_peak_sustained_active',
                        'smsp__inst_executed.sum', 'sm__inst_executed_pipe_', 'sm__pipe_fp64', 'smsp__issue_active.avg.pct',
                        'sm__throughput.avg.pct', 'issue_stalled', 'l1tex__t_sector_hit_rate', 'lts__t_sector_hit_rate',
                        'sm__cycles_elapsed.avg', 'smsp__thread_inst_executed_per_inst', 'smsp__sass_thread_inst_executed_op_d',
                        'gpc__cycles_elapsed.max', 'sm__cycles_active.avg', 'l1tex__data_pipe_lsu_wavefronts', 'smsp__inst_issued']
for vals in rows[2:]:
    for h, u, v in zip(hdr, units, vals):
        if any(w in h for w in want):
            print('%-105s %-14s %s' % (h, u, v))
    print('-' * 20)
