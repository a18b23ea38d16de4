"""Per-kernel launch counts and durations from an `ncu --metrics gpu__time_duration.sum --csv --log-file F` launch list.
usage: python scripts/ncu_kernel_times.py F [name filter]"""
import collections
import csv
import sys


def main(path, flt=None):
    rows = list(csv.reader(open(path)))
    i = [k for k, r in enumerate(rows) if r and r[0] == 'ID'][0]
    h = rows[i]
    kn, mv, gs, bs = h.index('Kernel Name'), h.index('Metric Value'), h.index('Grid Size'), h.index('Block Size')
    d = collections.OrderedDict()
    for r in rows[i + 1:]:
        if len(r) > mv:
            k = r[kn].split('(')[0]
            if flt and flt not in k:
                continue
            e = d.setdefault(k, [0, 0.0, 1e30, r[gs], r[bs]])
            v = float(r[mv].replace(',', ''))
            e[0] += 1
            e[1] += v
            e[2] = min(e[2], v)
    tot = sum(e[1] for e in d.values())
    for k, e in sorted(d.items(), key=lambda x: -x[1][1]):
        print('%-60s n=%4d avg %10.1f us min %10.1f us share %.3f grid %s block %s' % (k[:60], e[0], e[1] / e[0] / 1e3, e[2] / 1e3, e[1] / tot, e[3], e[4]))


if __name__ == '__main__':
    main(*sys.argv[1:3])
