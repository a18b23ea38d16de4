"""CPU: lbg_scipy_cubic_eval_host (csrc/ctx.cu, host code of liblbgcmb.so) reproduces scipy's `interp1d(kind='cubic')` -- the
routine the reference evaluates its C_l with (prep_camb.py:88-89) -- BIT FOR BIT: the not-a-knot coefficients in LAPACK dgbsv's
operation order and the de Boor evaluation.  That is what lets the library alone (no scipy at run time) reach the reference to
1e-10 on the N0 rows below the first non-zero C_l, where 1/C^tot is pure detector noise and the reference's value hangs on the
last bits of this routine (DESIGN.md "parity notes").  No GPU needed."""
import ctypes as C

import numpy as np
from scipy.interpolate import interp1d, make_interp_spline


def _eval(y, x):
    from lbgcmb_b200 import _ffi
    y, x = _ffi.f64(y), _ffi.f64(x)
    out = np.empty(x.size)
    _ffi.check(_ffi.lib().lbg_scipy_cubic_eval_host(y.size, _ffi._dp(y), x.size, _ffi._dp(x), _ffi._dp(out)))
    return out


def test_bit_identical_to_interp1d_on_cmb_tables():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    ell, lens, nolens = cmb_cl_tables(5300)
    rng = np.random.default_rng(11)
    phi = np.arange(0.0, 2. * np.pi + 2. * np.pi / 200, 2. * np.pi / 200)
    one = ((1.0 * np.cos(phi)) ** 2. + (1.0 * np.sin(phi)) ** 2.) ** 0.5            # |u|_j of the l1 = 1 row: 1 +- an ulp
    x = np.concatenate([one, 1.5 * one, rng.uniform(0., 5299., 4000), np.arange(0., 5300., 7.), [-0.5, 5299.0, 5299.5]])
    for tab in (lens[0], lens[1], lens[2], lens[3], nolens[0], nolens[3]):
        ref = interp1d(ell, tab, kind='cubic', bounds_error=False, fill_value=0.0)(x)
        got = _eval(tab, x)
        assert np.array_equal(got, ref)
    # the values that matter: on the l1 = 1 row the spline is rounding noise around C_1 = 0, not zero
    v = _eval(lens[0], one)
    assert np.any(v != 0.0) and np.max(np.abs(v)) < 1e-24


def test_coefficients_bit_identical_on_rough_data():
    """Random (non-smooth) ordinates exercise the pivoting of the banded LU: still the same bits as make_interp_spline."""
    rng = np.random.default_rng(3)
    for n in (6, 7, 50, 1001):
        y = rng.standard_normal(n) * 10. ** rng.uniform(-12, 0, n)
        x = rng.uniform(0, n - 1, 500)
        ref = make_interp_spline(np.arange(n, dtype=np.float64), y, k=3)(x)
        assert np.array_equal(_eval(y, x), ref), n
