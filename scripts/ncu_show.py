"""Print metric name / value pairs of an `ncu --csv --metrics ...` log: python scripts/ncu_show.py file.csv [nodes]"""
import csv
import sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
im, iv, iu = hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
vals = {}
for r in rows[1:]:
    vals[r[im]] = float(r[iv].replace(',', ''))
    print('%-95s %-8s %s' % (r[im], r[iu], r[iv]))
if len(sys.argv) > 2:
    wn = float(sys.argv[2]) / 32.0
    print('--- per warp-node (nodes = %s)' % sys.argv[2])
    for k in ('smsp__inst_executed.sum', 'sm__inst_executed_pipe_fp64.sum', 'sm__inst_executed_pipe_alu.sum', 'sm__inst_executed_pipe_lsu.sum',
              'sm__inst_executed_pipe_xu.sum', 'sm__inst_executed_pipe_fma.sum', 'sm__inst_executed_pipe_uniform.sum', 'sm__inst_executed_pipe_cbu.sum',
              'sm__inst_executed_pipe_adu.sum', 'l1tex__data_pipe_lsu_wavefronts.sum', 'l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum',
              'l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum'):
        if k in vals:
            print('%-60s %.1f' % (k, vals[k] / wn))
    if 'sm__cycles_elapsed.avg' in vals:
        print('%-60s %.1f' % ('SMSP-clk per warp-node', vals['sm__cycles_elapsed.avg'] * 148 * 4 / wn))
