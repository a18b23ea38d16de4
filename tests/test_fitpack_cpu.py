"""CPU: the library's restatement of FITPACK's interpolating quintic spline (csrc/fitpack.cuh, through the host-only
lbg_halofit_spec_host) against scipy's InterpolatedUnivariateSpline(k=5), which is what modi/halofit.py:70-98 calls: R_nl from
the inverse fit, n_eff and the curvature from the first / second derivative of the forward fit -- BIT FOR BIT, also on the
reference's golden spectral parameters (tests/golden/halofit.npz `spec`, written by the unmodified `_get_spec`) when fed the
oracle's sigma^2 samples.  No GPU needed."""
import os

import numpy as np
from scipy.interpolate import InterpolatedUnivariateSpline as spline


def _scipy_spec(lnr, lnsig):
    r_of_sig = spline(lnsig[::-1], lnr[::-1], k=5)
    rknl = 1.0 / np.exp(r_of_sig(0.0))
    dev1, dev2 = spline(lnr, lnsig, k=5).derivatives(np.log(1.0 / rknl))[1:3]
    return np.array([rknl, -dev1 - 3.0, -dev2])


def test_spec_bit_identical_to_scipy():
    from lbgcmb_b200.modi.halofit import _spec_host
    lnr = np.linspace(np.log(0.1), np.log(10.0), 500)
    rng = np.random.default_rng(2)
    for shift in (0.0, -1.3, -3.1):
        lnsig = -1.7 * lnr - 0.12 * lnr ** 2 + 0.8 + 0.02 * np.sin(2.5 * lnr) + shift + 1e-13 * rng.standard_normal(lnr.size)
        got, status = _spec_host(lnr, lnsig)
        if lnsig[0] > 0 > lnsig[-1]:
            assert status == 0
            assert np.array_equal(got, _scipy_spec(lnr, lnsig)), shift
        else:
            assert status == 1                    # ln sigma^2 does not cross zero on the grid: the reference's except branch


def test_reference_spec_from_reference_sigma2(golden_dir):
    """sigma^2(R) summed exactly as modi/halofit.py:36-44 does (numpy + the scipy-1.2.1 simps rule) + the library's fits reproduce the
    UNMODIFIED reference's (k_nl, n_eff, curvature) goldens to the last bits: the fits add nothing to the sigma^2 rounding."""
    from oracle.simps121 import simps
    from lbgcmb_b200.modi.halofit import spec_from_lnsig
    g = np.load(os.path.join(golden_dir, 'halofit.npz'))
    k, p = g['k'], g['plin']
    lnr = np.linspace(np.log(0.1), np.log(10.0), 500)
    lnk = np.log(k)
    for i in (0, 2, 5):
        delta = p * k ** 3 / (2. * np.pi ** 2) * g['D'][i] ** 2.
        lnsig = np.array([np.log(simps(delta * np.exp(-(k * np.exp(r)) ** 2), lnk)) for r in lnr])
        got = np.array(spec_from_lnsig(lnr, lnsig))
        assert np.allclose(got, g['spec'][i], rtol=1e-14, atol=0), (i, got, g['spec'][i])


def test_quotient_from_rounded_reciprocal_is_correctly_rounded():
    """halofit_spec_kernel's back-substitution replaces store / a0 by q0 = RN(store y), q <- fma(fma(-a0, q, store), y, q) twice, with
    y = RN(1 / a0) (csrc/halofit.cu): the claim is that this IS RN(store / a0), the reference's division.  Checked here with libm's fma
    on random and adversarial operands (mantissas near all-ones / near a power of two / short: ties and exact quotients); the same
    loop over 4e8 pairs in C found no mismatch either."""
    import ctypes
    import ctypes.util
    libm = ctypes.CDLL(ctypes.util.find_library('m'))
    libm.fma.restype = ctypes.c_double
    libm.fma.argtypes = [ctypes.c_double] * 3
    fma = libm.fma
    rng = np.random.default_rng(11)
    n = 60000

    def mant(kind, size):
        m = rng.integers(0, 1 << 52, size, dtype=np.uint64)
        if kind == 1:
            m = (np.uint64((1 << 52) - 1) - rng.integers(0, 256, size, dtype=np.uint64))
        elif kind == 2:
            m = rng.integers(0, 256, size, dtype=np.uint64)
        elif kind == 3:
            m = rng.integers(0, 1 << 20, size, dtype=np.uint64) << np.uint64(32)
        return m

    bad = 0
    for ka in range(4):
        for kb in range(4):
            ea = rng.integers(-20, 20, n // 16).astype(np.uint64) + np.uint64(1023)
            eb = rng.integers(-20, 20, n // 16).astype(np.uint64) + np.uint64(1023)
            a = ((ea << np.uint64(52)) | mant(ka, n // 16)).view(np.float64) * rng.choice([-1., 1.], n // 16)
            b = ((eb << np.uint64(52)) | mant(kb, n // 16)).view(np.float64)
            y = 1.0 / b
            want = a / b
            for ai, bi, yi, wi in zip(a.tolist(), b.tolist(), y.tolist(), want.tolist()):
                q = ai * yi
                q = fma(fma(-bi, q, ai), yi, q)
                q = fma(fma(-bi, q, ai), yi, q)
                bad += q != wi
    assert bad == 0
