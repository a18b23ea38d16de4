"""Pickle-cache key of the Nkk mirror (lbgcmb_b200/Nkk.py; reference cmb/Nkk.py:36-46): the reference's file name ignores the C_l
tables and the exact L grid, so a stale file is silently served for another cosmology; the mirror appends a digest of both."""
import numpy as np


def test_cache_key_follows_the_spectra(tmp_path, monkeypatch):
    from lbgcmb_b200 import Nkk
    from lbgcmb_b200.prep_camb import ClInterp
    from lbgcmb_b200.synthetic import cmb_cl_tables
    monkeypatch.setenv('LBGCMB', str(tmp_path))
    ell, lens, nolens = cmb_cl_tables(3000)
    li = {x: ClInterp(ell, lens[i]) for i, x in enumerate(('TT', 'EE', 'BB', 'TE'))}
    ni = {x: ClInterp(ell, nolens[i]) for i, x in enumerate(('TT', 'EE', 'BB', 'TE'))}
    Ls = np.arange(10, 60, 5)
    args = (Ls, ['TT', 'EB'], 1.4, 7.0, False)
    a = Nkk._cache_path(*args, li, ni)
    assert a == Nkk._cache_path(*args, li, ni) and a.endswith('.p') and '_cl_' in a
    assert a.startswith(Nkk._cache_path(*args)[:-2])                       # the reference's name is the prefix
    li2 = dict(li)
    t = np.array(Nkk.table_of(li['TT']), dtype=np.float64)
    t[100] *= 1. + 1e-12
    li2['TT'] = ClInterp(ell, t)
    assert Nkk._cache_path(*args, li2, ni) != a                            # one changed knot -> another file
    Ls2 = Ls.copy()
    Ls2[3] += 1                                                            # same length, ends: the reference's name cannot tell
    assert Nkk._cache_path(Ls2, *args[1:]) == Nkk._cache_path(*args)
    assert Nkk._cache_path(Ls2, *args[1:], li, ni) != a
