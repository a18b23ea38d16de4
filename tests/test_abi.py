"""CPU: the C-ABI library loads and exports exactly what include/lbgcmb.h declares; fails loudly w/o GPU."""
import ctypes
import os
import re

import numpy as np
import pytest

from lbgcmb_b200 import _ffi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_symbols():
    src = open(os.path.join(ROOT, 'include', 'lbgcmb.h')).read()
    src = re.sub(r'/\*.*?\*/', '', src, flags=re.S)
    return sorted(set(re.findall(r'\b(lbg_[a-z0-9_]+)\s*\(', src)))


def test_library_exports_every_declared_symbol():
    lib = _ffi.lib()
    syms = header_symbols()
    assert len(syms) >= 10
    for s in syms:
        assert hasattr(lib, s), 'missing export %s' % s


def test_ffi_declares_every_compute_symbol():
    src = open(os.path.join(ROOT, 'lbgcmb_b200', '_ffi.py')).read()
    for s in header_symbols():
        assert s in src, '_ffi.py does not bind %s' % s


def test_version_and_error_string():
    lib = _ffi.lib()
    assert b'sm_100a' in lib.lbg_version()
    rc = lib.lbg_cubic_spline_coeffs(2, None, None)
    assert rc == -1
    assert b'lbg_cubic_spline_coeffs' in lib.lbg_last_error()


def test_field_fisher_rejects_null_context_before_any_device_work():
    lib = _ffi.lib()
    rc = lib.lbg_field_fisher(None, 1, 1, 1, None, None, None, 1, None, None, None, None)
    assert rc == -1
    assert b'lbg_field_fisher' in lib.lbg_last_error()


def test_no_cpu_fallback_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip('GPU present')
    with pytest.raises(_ffi.LbgError):
        _ffi.Context(0)


def test_spline_coefficients_match_scipy_interp1d(cl_tables):
    from scipy.interpolate import interp1d
    ell, lens, _ = cl_tables
    lib = _ffi.lib()
    n = ell.size
    rng = np.random.default_rng(3)
    x = rng.uniform(2, n - 1, 50000)
    x[:4] = [2.0, n - 1.0, n - 2.0, 2.5]
    i = np.minimum(x.astype(int), n - 2)
    t = x - i
    um = 1 - t
    w = [um ** 3 / 6, 2 / 3 - t * t + .5 * t ** 3, 2 / 3 - um * um + .5 * um ** 3, t ** 3 / 6]
    for k in range(3):   # TT, EE, BB are positive: relative comparison is meaningful
        y = _ffi.f64(lens[k])
        coef = np.zeros(n + 2)
        _ffi.check(lib.lbg_cubic_spline_coeffs(n, _ffi._dp(y), _ffi._dp(coef)))
        s = sum(w[m] * coef[i + m] for m in range(4))
        r = interp1d(ell, y, kind='cubic')(x)
        assert np.max(np.abs(s / r - 1)) < 1e-13
