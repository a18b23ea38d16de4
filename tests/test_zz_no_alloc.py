"""GPU: lbg_snr_noalloc (csrc/fisher.cu) -- the allocation-free variant of lbg_snr whose device scratch is a fixed file-scope array --
against lbg_snr itself through the raw C ABI, on two calls in a row (the second on different data, so stale scratch would show), at
the array's limit, and beyond it (LBG_EINVAL).  Collected last on purpose (file name).
Tolerance: the same kernel on the same inputs -> identical bits (asserted with the suite's 1e-10 as the outer bound)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10
EINVAL = -1


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def _case(n, seed):
    rng = np.random.default_rng(seed)
    ell = np.arange(50., 50. + n)
    cl = 1e-7 / ell * (1. + 0.1 * rng.standard_normal(n))
    var = (1e-7 / ell) ** 2 * (2. + rng.random(n)) / (2. * ell + 1.)
    nm = rng.integers(1, 9, n).astype(np.float64)
    return ell, cl, var, nm


def test_snr_noalloc_matches_snr_on_two_calls_in_a_row():
    from lbgcmb_b200 import _ffi
    L = _ffi.lib()
    h = _ffi.context(0).handle                       # an existing context: streams / events are created with it, not per call
    cases = [(_case(4951, 1), 3000.), (_case(700, 2), 1e5), (_case(16384, 3), 9000.)]    # second call: shorter, other data
    for (ell, cl, var, nm), lmax in cases:
        want, got = C.c_double(-1.), C.c_double(-2.)
        assert L.lbg_snr(h, ell.size, _dp(ell), _dp(cl), _dp(var), _dp(nm), lmax, C.byref(want)) == 0
        assert L.lbg_snr_noalloc(h, ell.size, _dp(ell), _dp(cl), _dp(var), _dp(nm), lmax, C.byref(got)) == 0, L.lbg_last_error()
        assert want.value > 0 and abs(got.value / want.value - 1) < RTOL
        assert got.value == want.value               # same kernel, same inputs: same bits
        ref = np.sqrt(np.sum((nm * cl ** 2 / var)[ell <= lmax]))
        assert abs(got.value / ref - 1) < RTOL
    # nothing below the cut -> 0 (the result slot does not keep the previous call's value)
    ell, cl, var, nm = cases[1][0]
    got = C.c_double(-2.)
    assert L.lbg_snr_noalloc(h, ell.size, _dp(ell), _dp(cl), _dp(var), _dp(nm), 10., C.byref(got)) == 0 and got.value == 0.0
    assert L.lbg_snr_noalloc(h, 0, None, None, None, None, 1e5, C.byref(got)) == 0 and got.value == 0.0


def test_snr_noalloc_rejects_sizes_beyond_its_scratch():
    from lbgcmb_b200 import _ffi
    L = _ffi.lib()
    h = _ffi.context(0).handle
    ell, cl, var, nm = _case(16385, 4)
    got = C.c_double(-2.)
    assert L.lbg_snr_noalloc(h, ell.size, _dp(ell), _dp(cl), _dp(var), _dp(nm), 1e5, C.byref(got)) == EINVAL
    assert b'16384' in L.lbg_last_error() and got.value == -2.
    assert L.lbg_snr_noalloc(None, 4, _dp(ell), _dp(cl), _dp(var), _dp(nm), 1e5, C.byref(got)) == EINVAL
