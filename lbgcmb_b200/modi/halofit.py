"""HALOFIT; mirrors reference modi/halofit.py:7-214 (`_get_spec`, `halofit`), itself a copy of hmf's.

Device work (csrc/pk_kernels.cuh): the sigma^2(R) Simpson integrals over the k table (`sigma2_kernel`, halofit.py:36-44)
and the Smith/Takahashi formula on every table node (`pnl_kernel`, halofit.py:139-212).  The quintic-spline glue
(k_nl, n_eff, curvature; halofit.py:70-98) is the library's restatement of the FITPACK routines behind the reference's
`InterpolatedUnivariateSpline(k=5)` (csrc/fitpack.cuh: same statements, same order, bit-identical coefficients and
derivatives) -- an analytic shortcut, or any other correct spline code, misses 1e-10 on the curvature (SURVEY.md "hard
parts").  For whole z grids the same fits run on the device (csrc/halofit.cu, lbg_pk_build_table_hf)."""
import ctypes as _C

import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200.quadrature import simps_weights_x


class _DeviceTable(object):
    """(k, Delta^2) uploaded as a linear P(k) table on the context; sigma^2(R) evaluator."""

    def __init__(self, k, delta_k, device=None):
        self.ctx = _ffi.context(device)
        self.k = _ffi.f64(k)
        p0 = _ffi.f64(delta_k) * (2. * np.pi ** 2) / self.k ** 3
        w = _ffi.f64(simps_weights_x(np.log(self.k)))
        _ffi.check(_ffi.lib().lbg_upload_pk_linear(self.ctx.handle, self.k.size, _ffi._dp(self.k), _ffi._dp(p0), _ffi._dp(w)))
        _ffi.table_owner[self.ctx.device] = (_ffi.new_owner_token(), None)      # a one-off table: nobody else's

    def sigma2(self, R):
        R = _ffi.f64(np.atleast_1d(R))
        out = np.empty(R.size)
        _ffi.check(_ffi.lib().lbg_halofit_sigma2(self.ctx.handle, R.size, _ffi._dp(R), _ffi._dp(out)))
        return out


def _spec_host(lnr, lnsig, rknl=0.0):
    """(R_nl, n_eff, curvature), status: the two quintic fits of modi/halofit.py:70-78 in FITPACK order (lbg_halofit_spec_host)."""
    lnr, lnsig = _ffi.f64(lnr), _ffi.f64(lnsig)
    out = np.empty(3)
    status = _C.c_int(0)
    _ffi.check(_ffi.lib().lbg_halofit_spec_host(lnr.size, _ffi._dp(lnr), _ffi._dp(lnsig), float(rknl), _ffi._dp(out), _C.byref(status)))
    return out, status.value


def spec_from_lnsig(lnr, lnsig, sigma2_of_R=None):
    """k_nl, n_eff, curvature from samples of ln sigma^2(ln R); modi/halofit.py:70-98."""
    out, status = _spec_host(lnr, lnsig)
    rknl = out[0]
    if status:                                                     # the reference's `except` branch (:79-96)
        if sigma2_of_R is None:
            raise ValueError('derivative point ln(1 / R_nl) outside the fitted range')
        print("WARNING: ", 'derivative point ln(1 / R_nl) = %g outside the fitted range' % np.log(1.0 / rknl))
        lnr = np.linspace(np.log(0.2 / rknl), np.log(5 / rknl), 100)
        lnsig = np.log(sigma2_of_R(np.exp(lnr)))
        keep = np.logical_not(np.isinf(lnsig))
        lnr, lnsig = lnr[keep], lnsig[keep]
        if len(lnr) < 2:
            raise Exception("Lots of things went wrong in halofit")
        out, status = _spec_host(lnr, lnsig, rknl=rknl)
        if status:
            raise Exception("Lots of things went wrong in halofit")
    return rknl, out[1], out[2]


def _get_spec_from(sigma2_of_R, sigma_8):
    if sigma_8 < 1.0 and sigma_8 > 0.6:
        lnr = np.linspace(np.log(0.1), np.log(10.0), 500)
        lnsig = np.log(sigma2_of_R(np.exp(lnr)))
    else:  # weird sigma_8: search the decade where ln sigma crosses zero (modi/halofit.py:46-68)
        lnsig_old = None
        r = None
        for r in [0.00001, 0.0001, 0.001, 0.01, 0.1, 1.0, 10.0, 100.0, 1000.0]:
            lnsig1 = np.log(sigma2_of_R(np.array([r]))[0])
            if lnsig1 < 0:
                if lnsig_old is None:
                    print("WARNING: LOWEST R NOT LOW ENOUGH IN _GET_SPEC. ln(sig) starts below 0: ", lnsig1)
                break
            lnsig_old = lnsig1
        lnr = np.linspace(np.log(0.1 * r), np.log(r), 250)
        lnsig = np.log(sigma2_of_R(np.exp(lnr)))
    return spec_from_lnsig(lnr, lnsig, sigma2_of_R)


def _get_spec(k, delta_k, sigma_8):
    """Non-linear wavenumber, effective spectral index and curvature (modi/halofit.py:7-100)."""
    tab = _DeviceTable(k, delta_k)
    return _get_spec_from(tab.sigma2, sigma_8)


def halofit(k, delta_k, sigma_8, z, cosmo=None, takahashi=True):
    """Non-linear dimensionless power at `k` (modi/halofit.py:102-214).  `cosmo` is duck-typed as in the reference:
    .Om(z), .Ode(z), .w(z), .Onu0, .Om0."""
    tab = _DeviceTable(k, delta_k)
    rknl, neff, rncur = _get_spec_from(tab.sigma2, sigma_8)
    fnu = cosmo.Onu0 / cosmo.Om0
    row = _ffi.f64([[1.0, rknl, neff, rncur, float(cosmo.Om(z)), float(cosmo.Ode(z)), 1.0, 0.0]])
    out = np.empty((1, tab.k.size))
    _ffi.check(_ffi.lib().lbg_pk_build_table(tab.ctx.handle, 1, _ffi._dp(row), float(fnu), float(cosmo.Om0),
                                             float(cosmo.w(z)), int(bool(takahashi)), _ffi._dp(out)))
    return out[0] * tab.k ** 3 / (2. * np.pi ** 2)
