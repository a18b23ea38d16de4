"""GPU parity: lbg_n0_modi (csrc/n0_modi.cu) through the mirror of modi/nkk.py vs golden vectors written by the
UNMODIFIED reference (tests/golden/n0_modi.npz: N_fisher / N_fisherTE on in-memory splines and the CMB class on
CLASS-format files) and the numpy oracle.  Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


@pytest.fixture(scope='module')
def g(golden_dir):
    return np.load(os.path.join(golden_dir, 'n0_modi.npz'))


def _tables():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    ell, lens, nolens = cmb_cl_tables(5801)
    eu, el = ell[2:5801], ell[2:5301]
    u = {'TT': nolens[0, 2:5801], 'EE': nolens[1, 2:5801], 'BB': nolens[2, 2:5801], 'TE': nolens[3, 2:5801]}
    le = {'TT': lens[0, 2:5301], 'EE': lens[1, 2:5301], 'BB': lens[2, 2:5301], 'TE': lens[3, 2:5301]}
    return eu, el, u, le


def test_spline_ext_matches_fitpack(ctx):
    from scipy.interpolate import InterpolatedUnivariateSpline as IUS
    from lbgcmb_b200.modi.nkk import interpolate
    eu, el, u, le = _tables()
    # inside the knots, plus the short extrapolations the reference actually makes (l in [1, 2) below the first CLASS
    # multipole); far extrapolation of a cubic amplifies last-bit coefficient differences and is not a parity target
    x = np.concatenate([np.random.default_rng(2).uniform(1.0, el[-1], 20000), [1.0, 1.5, 2.0, el[-1], el[-1] + 1.5]])
    for y in (le['TT'], le['EE'], u['TT']):
        xs = eu if y.size == eu.size else el
        got, ref = interpolate(xs, y)(x), IUS(xs, y)(x)
        assert np.max(np.abs(got / ref - 1)) < 1e-11
    with pytest.raises(ValueError):
        interpolate(np.array([2., 3., 5., 6.]), np.ones(4))


def test_n_fisher_golden(ctx, g):
    from lbgcmb_b200.DetectorNoise import DetectorNoise
    from lbgcmb_b200.modi import nkk
    eu, el, u, le = _tables()
    thetab, delT, Ls = float(g['thetab']), float(g['delT']), g['Ls']
    I = nkk.interpolate
    for mode in ('TT', 'EE', 'EB'):
        a1, a2 = mode[0] * 2, mode[1] * 2
        lint = np.linspace(1., 5000. if mode in ('EB', 'EE') else 3000., 10000)
        args = (I(eu, u[a1]), I(eu, u[a2]), I(el, le[a1] + DetectorNoise(el, thetab, delT, a1)), I(el, le[a2] + DetectorNoise(el, thetab, delT, a2)))
        got = nkk.N_fisher(Ls, lint, *args, mode)
        assert np.max(np.abs(got / g['int_' + mode] - 1)) < RTOL, mode
        one = nkk.N_fisher(float(Ls[1]), lint, *args, mode)          # scalar L, as the reference calls it
        assert isinstance(one, float) and one == got[1]
    lint = np.linspace(1., 3000., 10000)
    args = (I(eu, u['TE']), I(el, le['TT'] + DetectorNoise(el, thetab, delT, 'TT')), I(el, le['EE'] + DetectorNoise(el, thetab, delT, 'EE')), I(el, le['TE']))
    got = nkk.N_fisherTE(Ls, lint, *args)
    assert np.max(np.abs(got / g['int_TE'] - 1)) < RTOL
    with pytest.raises(ValueError):
        nkk.N_fisher(Ls, lint, *args, 'TE')


def test_n_fisher_accepts_scipy_splines_and_matches_oracle(ctx):
    from scipy.interpolate import InterpolatedUnivariateSpline as IUS
    from lbgcmb_b200.modi import nkk
    from oracle import n0 as on0
    eu, el, u, le = _tables()
    lint = np.linspace(1., 3000., 2000)          # a coarser l grid keeps the numpy oracle to ~1 s
    n1 = on0.detector_noise(el, 3.0, 10.0, 'TT', 2.786)
    args = (IUS(eu, u['TT']), IUS(eu, u['TT']), IUS(el, le['TT'] + n1), IUS(el, le['TT'] + n1))
    L = 431.7
    got, ref = nkk.N_fisher(L, lint, *args, 'TT'), on0.modi_N_fisher(L, lint, *args, 'TT')
    assert abs(got / ref - 1) < RTOL
    with pytest.raises(TypeError):
        nkk.N_fisher(L, lint, lambda x: x, args[1], args[2], args[3], 'TT')


def test_cmb_class_on_class_files(ctx, g, tmp_path):
    from lbgcmb_b200.modi import nkk
    from lbgcmb_b200.synthetic import write_class_files
    clf, clsf = str(tmp_path / 'cl.dat'), str(tmp_path / 'cl_lensed.dat')
    write_class_files(clf, clsf)
    cmb = nkk.CMB(clfile=clf, clsfile=clsf)
    assert np.array_equal(cmb.lat, g['lat'])
    thetab, delT = float(g['thetab']), float(g['delT'])
    Nl1, Nl2, ntt = cmb.do_noise('T', 'T', thetab, delT)
    assert np.max(np.abs(ntt / g['cmb_TT'] - 1)) < RTOL
    assert np.max(np.abs(cmb.do_noise('E', 'B', thetab, delT)[2] / g['cmb_EB'] - 1)) < RTOL
    assert np.max(np.abs(cmb.do_noiseTE(thetab, delT)[2] / g['cmb_TE'] - 1)) < RTOL
    assert cmb.do_noise('T', 'E', thetab, delT) is None                 # "Wrong function", modi/nkk.py:200-203
    assert np.allclose(Nl2, nkk.N_inst(thetab, delT, cmb.ells, 'T'), rtol=1e-12)


def test_n_fisher_extreme_L_against_oracle(ctx):
    """modi flavour at the ends: L below the first l node, L next to / beyond the last one (part or all of the l2 = |L - l| range
    falls outside [min l, max l] and `frac` is zeroed there, modi/nkk.py:84-85; beyond 2 max l the integral is exactly 0)."""
    from scipy.interpolate import InterpolatedUnivariateSpline as IUS
    from lbgcmb_b200.modi import nkk
    from oracle import n0 as on0
    eu, el, u, le = _tables()
    lint = np.linspace(1., 3000., 1500)
    nT = on0.detector_noise(el, 1.4, 7.0, 'TT', 2.786)
    nE = on0.detector_noise(el, 1.4, 7.0, 'EE', 2.786)
    Ls = np.array([1.5, 2.0, 2999.5, 3000.0, 3500.0, 5999.0, 7000.0])
    cases = {'TT': (IUS(eu, u['TT']), IUS(eu, u['TT']), IUS(el, le['TT'] + nT), IUS(el, le['TT'] + nT)),
             'EE': (IUS(eu, u['EE']), IUS(eu, u['EE']), IUS(el, le['EE'] + nE), IUS(el, le['EE'] + nE)),
             'EB': (IUS(eu, u['EE']), IUS(eu, u['BB']), IUS(el, le['EE'] + nE), IUS(el, le['BB'] + nE))}
    for mode, args in cases.items():
        got = nkk.N_fisher(Ls, lint, *args, mode)
        ref = np.array([on0.modi_N_fisher(float(L), lint, *args, mode) for L in Ls])
        assert got[-1] == 0.0 and ref[-1] == 0.0
        ok = ref != 0
        assert np.max(np.abs(got[ok] / ref[ok] - 1)) < RTOL, (mode, got / ref - 1)
    args = (IUS(eu, u['TE']), IUS(el, le['TT'] + nT), IUS(el, le['EE'] + nE), IUS(el, le['TE']))
    got = nkk.N_fisherTE(Ls, lint, *args)
    ref = np.array([on0.modi_N_fisherTE(float(L), lint, *args) for L in Ls])
    ok = ref != 0
    assert np.array_equal(got == 0, ref == 0) and np.max(np.abs(got[ok] / ref[ok] - 1)) < RTOL, got / ref - 1
