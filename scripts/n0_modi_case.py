"""Timing case for lbg_n0_modi (csrc/n0_modi.cu n0_modi_kernel): the reference's own sizes -- modi/nkk.py CMB.do_noise / do_noiseTE,
15 L values x 10000 l nodes x 200 phi nodes per estimator -- on the synthetic CLASS-format tables.  One JSON line per estimator:
device milliseconds (CUDA events around the call's kernels), nodes per second.  Usage: python scripts/n0_modi_case.py [reps]"""
import json
import os
import sys
import tempfile

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from lbgcmb_b200 import synthetic                                   # noqa: E402
from lbgcmb_b200.modi import nkk                                    # noqa: E402


def main():
    reps = int(sys.argv[1]) if len(sys.argv) > 1 else 5
    d = tempfile.mkdtemp()
    clfile, clsfile = os.path.join(d, 'cl.dat'), os.path.join(d, 'cl_lensed.dat')
    synthetic.write_class_files(clfile, clsfile)
    cmb = nkk.CMB(clfile, clsfile)
    cases = [('TT', lambda: cmb.do_noise('T', 'T', 1.4, 7.0)), ('EE', lambda: cmb.do_noise('E', 'E', 1.4, 7.0)),
             ('EB', lambda: cmb.do_noise('E', 'B', 1.4, 7.0)), ('TE', lambda: cmb.do_noiseTE(1.4, 7.0))]
    devnull = open(os.devnull, 'w')
    for name, fn in cases:
        ms = []
        for _ in range(reps + 1):
            so, sys.stdout = sys.stdout, devnull                      # the mirror prints the reference's progress lines
            try:
                out = fn()[2]
            finally:
                sys.stdout = so
            ms.append(nkk.last_stats['kernel_ms'])
        ms = np.array(ms[1:])
        nodes = nkk.last_stats['nodes']
        print(json.dumps({'estimator': name, 'n_L': 15, 'nodes': nodes, 'kernel_ms_median': float(np.median(ms)), 'kernel_ms_min': float(ms.min()),
                          'launches': nkk.last_stats['launches'], 'gnodes_per_s': nodes / np.median(ms) / 1e6,
                          'finite': bool(np.all(np.isfinite(out)))}))


if __name__ == '__main__':
    main()
