"""Fraction of the folded (L, l1, phi) quadrature nodes of the bench workload that survive the |v| masks of
cmb/Nkk.py:13-26 (2 <= |v| <= 5000): the nodes the rows5 kernel actually visits (per lane; a warp walks the union of its
lanes' ranges).  python scripts/count_unmasked_nodes.py [lmin lmax]"""
import sys
import numpy as np

lmin, lmax = (int(sys.argv[1]), int(sys.argv[2])) if len(sys.argv) > 2 else (50, 5000)
l1 = np.arange(1.0, 5000.5, 0.5)
phi = np.arange(0.0, 2. * np.pi + 2. * np.pi / 200, 2. * np.pi / 200)[:101]
c = np.cos(phi)[::-1]                       # ascending for searchsorted
tot = kept = keptT = 0
for L in range(lmin, lmax + 1):
    # |v|^2 = l^2 + L^2 - 2 l L cos  <= 5000^2   <=>  cos >= (l^2 + L^2 - 5000^2) / (2 l L)
    for cut, sel in ((5000., None), (3000., l1 < 3000.)):
        x = (l1 ** 2 + L ** 2 - cut ** 2) / (2. * l1 * L)
        n_ok = 101 - np.searchsorted(c, x, side='left')
        xlo = (l1 ** 2 + L ** 2 - 4.) / (2. * l1 * L)           # |v| >= 2
        n_ok = n_ok - (101 - np.searchsorted(c, xlo, side='right'))
        n_ok = np.clip(n_ok, 0, 101)
        if sel is None:
            kept += n_ok.sum()
        else:
            keptT += n_ok[sel].sum()
    tot += 101 * l1.size
print('folded nodes %d, inside the polarisation masks %d (%.4f), inside the temperature masks %d (%.4f)'
      % (tot, kept, kept / tot, keptT, keptT / tot))
