"""CPU: the host-only mirrors of SURVEY.md 8(a) rows a8 (`prep_lmaxmask`) and a21 (`plot_ellipse`) against golden vectors
written by the UNMODIFIED reference (tests/golden/variances.npz, oracle/make_golden.py make_variances), and the variance
arithmetic of rows a3 / a5 / a6 (Cgg.py:70-83, Ckg.py:60-87, cmb/lensing.py:54-66) restated in oracle/limber.py."""
import os

import numpy as np
import pytest


@pytest.fixture(scope='module')
def g(golden_dir):
    return np.load(os.path.join(golden_dir, 'variances.npz'))


def test_prep_lmaxmask_golden(g):
    from lbgcmb_b200.Nkk import prep_lmaxmask
    m = prep_lmaxmask(g['mask_nu'], g['mask_nv'])
    assert sorted(m) == [str(k) for k in g['mask_keys']]
    for key, ref in zip(g['mask_keys'], g['masks']):
        assert np.array_equal(m[str(key)], ref), key
    assert m['TT'] is m['TE'] and m['EE'] is m['EB']                 # the reference hands out two arrays, not six


def test_plot_ellipse_golden(g):
    """fisher_contour.py:64-111 with cov= and data_out=True: SVD axes, chi^2 table lookup, rotation, centre."""
    from lbgcmb_b200.fisher_contour import plot_ellipse, ellipse_from_cov
    from oracle.make_golden import variance_inputs
    covs = variance_inputs()['covs']
    for i, cov in enumerate(covs):
        for tag, ml in (('68', 0.68), ('95', 0.95), ('none', None)):
            got = plot_ellipse(cov=cov.copy(), mass_level=ml, data_out=True, theta_num=1000, x_cent=0.8, y_cent=3.9)
            ref = g['ell_%d_%s' % (i, tag)]
            assert got.shape == ref.shape == (2, 1000)
            assert np.max(np.abs(got - ref)) <= 1e-13 * np.max(np.abs(ref)), (i, tag)
    got = plot_ellipse(semimaj=2.0, semimin=0.5, phi=0.3, theta_num=500, data_out=True)
    assert np.max(np.abs(got - g['ell_axes'])) < 1e-14
    assert plot_ellipse(cov=covs[0], data_out=True).shape == (2, 1000)        # the default theta_num = 1e3 is a float
    a, b, ang = ellipse_from_cov(covs[1], 0.68)
    assert a > b > 0 and -np.pi / 2 <= ang <= np.pi / 2
    # 68 % of a 2-d Gaussian lies inside chi^2 = 2.279 (the reference's mass_level=None constant): the two agree to the table step
    a0, _, _ = ellipse_from_cov(covs[1], None)
    assert abs(a / a0 - 1) < 1e-3


def test_variance_arithmetic_oracle(g):
    """oracle/limber.py var_* on the golden spectra: the same three lines of arithmetic as the reference, bit for bit."""
    from oracle import limber as olim
    L, nkk = g['Llls'].astype(np.float64), g['nkk']
    ngg = olim.Ngg(L, 0.96, 4.96, lambda z: np.exp(-(z - 2.96) ** 2 / (2. * 0.24 ** 2)) / (np.sqrt(2. * np.pi) * 0.24), 2181.0)
    assert np.array_equal(olim.var_auto(L, g['cgg_full'], ngg, 0.4), g['var_cgg_given'])
