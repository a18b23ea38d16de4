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


def test_tables_handed_over_without_a_copy_when_they_already_form_a_matrix():
    """Nkk._cl_tables: four tables that are the rows of one C-contiguous [4, n] array (prep_camb.cl_interps_from_tables) are passed
    as that array; any other layout (separate arrays, permuted rows, a longer unlensed table) is stacked -- same values either way."""
    from lbgcmb_b200 import Nkk
    from lbgcmb_b200.prep_camb import ClInterp, cl_interps_from_tables
    from lbgcmb_b200.synthetic import cmb_cl_tables
    ell, lens, nolens = cmb_cl_tables(2000)
    order = ('TT', 'EE', 'BB', 'TE')
    li, ni = cl_interps_from_tables(ell, lens, nolens)
    n, lt, ut = Nkk._cl_tables(li, ni)
    assert n == 2000 and np.shares_memory(lt, lens) and np.array_equal(lt, lens) and np.array_equal(ut, nolens)
    sep = {x: ClInterp(ell, lens[i].copy()) for i, x in enumerate(order)}
    perm = {x: ClInterp(ell, lens[(i + 1) % 4]) for i, x in enumerate(order)}
    for tabs, want in ((sep, lens), (perm, lens[[1, 2, 3, 0]])):
        _, got, _ = Nkk._cl_tables(tabs, ni)
        assert not np.shares_memory(got, lens) and np.array_equal(got, want) and got.flags.c_contiguous
    short = {x: ClInterp(ell[:1500], lens[i, :1500]) for i, x in enumerate(order)}      # lensed table shorter than the unlensed one
    n2, lt2, ut2 = Nkk._cl_tables(short, ni)
    assert n2 == 1500 and np.array_equal(lt2, lens[:, :1500]) and np.array_equal(ut2, nolens[:, :1500]) and ut2.flags.c_contiguous
