"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE.  usage:  python -m oracle.make_golden [n0|limber|halofit|modi|fisher ...]
Each fixture stores the exact inputs (tables, grids, cos/sin of the phi nodes as computed by this
machine's numpy) next to the reference's outputs, so the replay on the GPU box needs neither
/root/reference nor bit-identical libm.
"""
import os
import sys
import time

import numpy as np

from . import refload, n0 as on0

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def make_n0():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    Nkk = refload.load('Nkk')                      # /root/reference/cmb/Nkk.py
    ell, lens, nolens = cmb_cl_tables(5300)
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    _, pphi, _, _ = on0.root_grids()
    cases = [
        ('n0_root_cmbs4', np.array([50, 3000]), ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True),
        ('n0_root_so_ttee', np.array([1000]), ['TT', 'EE'], 1.4, 7.0, False),
        ('n0_root_planck_eb', np.array([4999]), ['EB', 'TE'], 7.0, 30.0, False),
    ]
    for name, Ls, terms, thetab, DeltaT, it in cases:
        t = time.time()
        ref = Nkk.Nkk(li, ni, Ls, terms, thetab, DeltaT, iterative=it, pickleit=False)
        dt = time.time() - t
        ora = on0.nkk_root(li, ni, Ls, tuple(terms), thetab, DeltaT, it)
        print(name, 'reference %.1f s' % dt, ref, 'oracle rel diff', np.max(np.abs(ora / ref - 1)))
        np.savez_compressed(os.path.join(GOLD, name + '.npz'), ell=ell, lensed=lens, unlensed=nolens, Ls=Ls,
                            terms=np.array(terms), thetab=thetab, DeltaT=DeltaT, iterative=it, Tcmb=2.786,
                            cos_phi=np.cos(pphi), sin_phi=np.sin(pphi), nkk_reference=ref, ref_seconds=dt)


MAKERS = {'n0': make_n0}

if __name__ == '__main__':
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or sorted(MAKERS)
    for w in which:
        MAKERS[w]()
