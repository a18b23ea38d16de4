"""GPU parity of the multi-level N0 call (lbg_n0_ho02_levels, csrc/n0.cu n0_rows5m_kernel: several noise levels of one beam per pass
of the integrand) against the single-level call (itself pinned to the unmodified reference's goldens, tests/test_n0_gpu.py) and the
numpy oracle.  Tolerance 1e-10 against the oracle; 1e-12 against the single-level kernel (same arithmetic except that the noise is
added after the Horner sums instead of inside the coefficients)."""
import time

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def interps(cl_tables):
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    return cl_interps_from_tables(*cl_tables)


def test_levels_match_single_level_calls_and_oracle(ctx, cl_tables, interps):
    from lbgcmb_b200 import Nkk
    from oracle import n0 as on0
    li, ni = interps
    ell, lens, nolens = cl_tables
    Ls = np.array([50, 333, 1500, 2999, 3001, 4000, 5000])
    dts = [1.0, 3.0, 7.0, 12.0, 30.0, 2.0, 0.5]
    its = [1, 0, 0, 1, 0, 0, 1]
    got, terms = Nkk.n0_levels(li, ni, Ls, 1.4, dts, its)
    assert got.shape == (7, Ls.size) and Nkk.last_stats['path'] == 6
    for v, (dt, it) in enumerate(zip(dts, its)):
        one, t1 = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, dt, bool(it))
        assert np.max(np.abs(got[v] / one - 1)) < 1e-12, (v, np.abs(got[v] / one - 1))
        for key in ('TT', 'TE', 'EE', 'EB'):
            assert np.max(np.abs(terms[key][v] / t1[key] - 1)) < 1e-12, (v, key)
    ref = on0.nkk_root(on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens), Ls[[0, 3]], ('TT', 'TE', 'EE', 'EB'), 1.4, 12.0, True)
    assert np.max(np.abs(got[3][[0, 3]] / ref - 1)) < 1e-10


def test_level_result_does_not_depend_on_its_group(ctx, interps):
    """Groups of 3, 2 and 1 levels run different template instances of the kernel: a level's bits are the same in each."""
    from lbgcmb_b200 import Nkk
    li, ni = interps
    Ls = np.arange(50, 650, 7)
    dts = [1.0, 5.0, 9.0, 20.0]
    a = Nkk.n0_levels(li, ni, Ls, 3.0, dts)[0]                      # groups (1, 5, 9) + (20)
    b = Nkk.n0_levels(li, ni, Ls, 3.0, dts[::-1])[0][::-1]          # groups (20, 9, 5) + (1)
    c = np.array([Nkk.n0_levels(li, ni, Ls, 3.0, [d])[0][0] for d in dts])
    d2 = Nkk.n0_levels(li, ni, Ls, 3.0, dts[:2])[0]
    assert np.array_equal(a, b) and np.array_equal(a, c) and np.array_equal(a[:2], d2)
    assert np.all(np.diff(a, axis=0) > 0)                           # more noise, more reconstruction noise
    assert Nkk.n0_levels(li, ni, np.array([]), 3.0, dts)[0].shape == (4, 0)


def test_levels_reject_what_the_kernel_cannot_serve(ctx, interps):
    from lbgcmb_b200 import Nkk
    li, ni = interps
    with pytest.raises(ValueError):
        Nkk.n0_levels(li, ni, np.array([100.]), 10.0, [1.0, 2.0])    # beam beyond the folded-noise bound
    with pytest.raises(ValueError):
        Nkk.n0_levels(li, ni, np.array([100.]), 1.4, [1.0, -2.0])
    assert np.all(Nkk.n0_terms(li, ni, np.array([100.]), ('TT', 'TE', 'EE', 'EB'), 1.4, 1.0, False)[0] > 0)   # the context is usable


def test_levels_time_per_level(ctx, interps):
    """Device time per (L, level) against the single-level kernel on a 1000-L grid: recorded, and required not to be slower."""
    from lbgcmb_b200 import Nkk
    li, ni = interps
    Ls = np.arange(50, 5001, 5)
    Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 1.0, False)
    single = Nkk.last_stats['main_kernel_ms']
    dts = np.linspace(1., 30., 6)
    Nkk.n0_levels(li, ni, Ls, 1.4, dts)
    Nkk.n0_levels(li, ni, Ls, 1.4, dts)
    multi = Nkk.last_stats['main_kernel_ms'] / dts.size
    print('N0 integrand ms per level on %d L: single-level kernel %.3f, multi-level kernel %.3f (x%.2f)' % (Ls.size, single, multi, single / multi))
    assert multi < single
