"""C-ABI error behaviour on a live context (include/lbgcmb.h: 0 on success, negative LBG_E* otherwise, message in lbg_last_error):
argument checks, state checks (table not built yet), the table slots and the second-context constructor."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
EINVAL, ESTATE = -1, -3


def _dp(a):
    return a.ctypes.data_as(C.POINTER(C.c_double))


def test_error_codes_and_messages():
    from lbgcmb_b200 import _ffi
    L = _ffi.lib()
    h = C.c_void_p()
    assert L.lbg_ctx_create_lane(0, -1, C.byref(h)) == EINVAL and b'lane' in L.lbg_last_error()
    assert L.lbg_ctx_create_lane(10 ** 6, 0, C.byref(h)) == EINVAL and not h
    assert L.lbg_ctx_create_lane(0, 2, C.byref(h)) == 0 and h            # a further lane: own stream / scratch
    one = np.ones(8)
    rows = C.c_int(-1)
    # nothing uploaded yet on this fresh context
    assert L.lbg_halofit_sigma2(h, 8, _dp(one), _dp(one.copy())) == ESTATE
    assert L.lbg_pk_build_table(h, 1, _dp(one), 0.0, 0.3, -1.0, 1, None) == ESTATE
    prow = np.zeros(8, dtype=np.int32)
    assert L.lbg_limber(h, 1, 8, _dp(one), 8, _dp(one), _dp(one), prow.ctypes.data_as(C.POINTER(C.c_int)), 0, _dp(one.copy())) == ESTATE
    assert b'lbg_pk_build_table first' in L.lbg_last_error()
    assert L.lbg_pk_eval(h, 0, 8, _dp(one), 0, _dp(one.copy())) == ESTATE
    assert L.lbg_eval_cl(h, 0, 0, 8, _dp(one), _dp(one.copy())) == EINVAL              # no C_l tables uploaded
    # argument checks
    assert L.lbg_upload_pk_linear(h, 3, _dp(one), _dp(one), _dp(one)) == EINVAL         # n_k < 4
    k = np.array([1., 2., 2., 3., 4.])
    assert L.lbg_upload_pk_linear(h, 5, _dp(k), _dp(k), _dp(k)) == EINVAL and b'increasing' in L.lbg_last_error()
    assert L.lbg_upload_cl_tables(h, 3, _dp(one), _dp(one)) == EINVAL
    assert L.lbg_snr(h, 0, None, None, None, None, 1.0, None) != 0
    assert L.lbg_pk_select_slot(h, 4, C.byref(rows)) == EINVAL and L.lbg_pk_select_slot(h, -1, None) == EINVAL
    assert L.lbg_growth_table(h, 1, _dp(np.array([-0.3])), _dp(one), 8, _dp(np.linspace(-3., 0., 8)), _dp(one.copy()), None) == EINVAL
    assert L.lbg_growth_table(h, 1, _dp(one), _dp(one), 3, _dp(np.array([0., -1., -2.])), _dp(one.copy()), None) == EINVAL
    # slots: a table built in slot 1 is still there after visiting slot 2; a new k grid drops them all
    k = np.logspace(-3, 1, 64)
    w = np.gradient(np.log(k))
    assert L.lbg_upload_pk_linear(h, 64, _dp(k), _dp(1e3 * k ** -1.5), _dp(w)) == 0
    assert L.lbg_pk_select_slot(h, 1, C.byref(rows)) == 0 and rows.value == 0
    halo = np.zeros((3, 8))
    halo[:, 0] = [1.0, 0.5, 0.25]
    assert L.lbg_pk_build_table(h, 3, _dp(halo), 0.0, 0.3, -1.0, 1, None) == 0
    assert L.lbg_pk_select_slot(h, 2, C.byref(rows)) == 0 and rows.value == 0
    assert L.lbg_pk_select_slot(h, 1, C.byref(rows)) == 0 and rows.value == 3
    x, out = np.array([0.1, 1.0]), np.zeros(2)
    assert L.lbg_pk_eval(h, 1, 2, _dp(x), 0, _dp(out)) == 0
    assert np.allclose(out, 0.5 * np.interp(x, k, 1e3 * k ** -1.5), rtol=1e-13)      # linear row: P_lin x D^2
    assert L.lbg_pk_eval(h, 3, 2, _dp(x), 0, _dp(out)) == EINVAL                      # row beyond the table
    assert L.lbg_upload_pk_linear(h, 64, _dp(k), _dp(1e3 * k ** -1.5), _dp(w)) == 0
    assert L.lbg_pk_select_slot(h, 1, C.byref(rows)) == 0 and rows.value == 0
    assert L.lbg_ctx_destroy(h) == 0
