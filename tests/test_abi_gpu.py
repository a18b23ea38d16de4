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


def test_n0_c_only_reaches_reference_goldens(golden_dir):
    """Raw C-ABI calls, n_exact = 0, no scipy anywhere: lbg_upload_cl_tables + lbg_n0_ho02 reproduce the UNMODIFIED reference's N0
    (tests/golden/n0_root_*.npz, incl. L = 50 where the l1 < 2 rows carry up to 20 % of the TT integral) to 1e-10 -- the library
    evaluates those rows' u-side splines itself in scipy's operation order (lbg_n0_set_exact_below, default 2).  With that
    switched off the same call is off by ~1e-7 at L = 50: the rows matter."""
    import glob
    import os
    from lbgcmb_b200 import _ffi
    from lbgcmb_b200.Nkk import root_grids, TERM_BITS
    L = _ffi.lib()
    h = C.c_void_p()
    assert L.lbg_ctx_create_lane(0, 3, C.byref(h)) == 0
    ell, w_ell, pphi, w_phi = (np.ascontiguousarray(a) for a in root_grids())
    worst_off = 0.0
    for f in sorted(glob.glob(os.path.join(golden_dir, 'n0_root_*.npz'))):
        g = np.load(f)
        lens, nolens = np.ascontiguousarray(g['lensed']), np.ascontiguousarray(g['unlensed'])
        assert L.lbg_upload_cl_tables(h, lens.shape[1], _dp(lens), _dp(nolens)) == 0
        Ls = np.ascontiguousarray(g['Ls'], dtype=np.float64)
        mask = 0
        for t in g['terms']:
            mask |= TERM_BITS[str(t)]
        cphi, sphi = np.ascontiguousarray(g['cos_phi']), np.ascontiguousarray(g['sin_phi'])
        res = {}
        for below in (2.0, 0.0):
            assert L.lbg_n0_set_exact_below(h, below) == 0
            out = np.empty(Ls.size)
            rc = L.lbg_n0_ho02(h, Ls.size, _dp(Ls), mask, float(g['thetab']), float(g['DeltaT']), float(g['Tcmb']), int(bool(g['iterative'])),
                               ell.size, _dp(ell), _dp(w_ell), pphi.size, _dp(cphi), _dp(sphi), _dp(w_phi), 3000., 5000., 2., 1,
                               0, None, None, _dp(out), None, None)
            assert rc == 0, L.lbg_last_error()
            res[below] = out
        assert np.max(np.abs(res[2.0] / g['nkk_reference'] - 1)) < 1e-10, os.path.basename(f)
        worst_off = max(worst_off, float(np.max(np.abs(res[0.0] / g['nkk_reference'] - 1))))
    assert 1e-10 < worst_off < 1e-5          # without the scipy-order rows the library is visibly (but only slightly) off
    assert L.lbg_ctx_destroy(h) == 0


def test_n0_begin_end_state_machine(cl_tables):
    from lbgcmb_b200 import _ffi
    from lbgcmb_b200.Nkk import root_grids
    L = _ffi.lib()
    h = C.c_void_p()
    assert L.lbg_ctx_create_lane(0, 3, C.byref(h)) == 0
    ell_t, lens, nolens = cl_tables
    ell, w_ell, pphi, w_phi = (np.ascontiguousarray(a) for a in root_grids())
    cphi, sphi = np.cos(pphi), np.sin(pphi)
    Ls = np.array([100., 2000.])
    out = np.empty(2)
    path = C.c_int(0)

    def begin(n_L=2, fold=1, lmin_v=2.):
        return L.lbg_n0_ho02_begin(h, n_L, _dp(Ls), 15, 1.4, 7.0, 2.786, 0, ell.size, _dp(ell), _dp(w_ell), pphi.size, _dp(cphi), _dp(sphi),
                                   _dp(w_phi), 3000., 5000., lmin_v, fold, 0, None, None)
    assert begin() == ESTATE and b'lbg_upload_cl_tables' in L.lbg_last_error()            # no tables yet
    assert L.lbg_n0_ho02_end(h, 2, _dp(out), None, None, None) == ESTATE                   # nothing pending
    lens, nolens = np.ascontiguousarray(lens), np.ascontiguousarray(nolens)
    assert L.lbg_upload_cl_tables(h, lens.shape[1], _dp(lens), _dp(nolens)) == 0
    assert begin(fold=3) == EINVAL and begin(lmin_v=0.0) == EINVAL and b'lmin_v' in L.lbg_last_error()
    assert begin() == 0
    assert begin() == ESTATE and b'pending' in L.lbg_last_error()                          # one call in flight per context
    assert L.lbg_upload_cl_tables(h, lens.shape[1], _dp(lens), _dp(nolens)) == ESTATE      # tables are in use
    assert L.lbg_n0_ho02_end(h, 3, _dp(out), None, None, None) == EINVAL                   # wrong size: the call is dropped
    assert begin() == 0
    unmasked = C.c_longlong(0)
    assert L.lbg_n0_ho02_end(h, 2, _dp(out), None, None, C.byref(unmasked)) == 0
    assert np.all(out > 0) and unmasked.value > 0
    assert L.lbg_n0_last_path(h, C.byref(path)) == 0 and path.value == 5
    assert begin(n_L=0) == 0 and L.lbg_n0_ho02_end(h, 0, None, None, None, None) == 0      # empty call
    assert L.lbg_ctx_destroy(h) == 0
