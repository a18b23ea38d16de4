"""Eisenstein & Hu (1999) transfer function and the linear P(k) table built from it; mirrors the C API of the reference's
EisensteinHu/EisensteinAndHu.c (`TFmdm_set_cosm` :101, `TFmdm_onek_mpc` :221, `TFmdm_onek_hmpc` :268) with arrays in place of
one wavenumber per call and an object in place of its global state.  The fit itself runs on the GPU (csrc/eh.cu)."""
import ctypes as C

import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200.quadrature import simps_weights_x


class TFmdm(object):
    """`TFmdm_set_cosm(omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble, redshift)` as an object."""

    def __init__(self, omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble, redshift, device=None):
        if hubble <= 0.0:
            raise ValueError('TFmdm_set_cosm(): Negative Hubble constant illegal.')          # the reference exit(1)s here (:135-138)
        if redshift <= -1.0:
            raise ValueError('TFmdm_set_cosm(): Redshift < -1 is illegal.')                   # (:145-148)
        self.args = (float(omega_matter), float(omega_baryon), float(omega_hdm), int(degen_hdm), float(omega_lambda),
                     float(hubble), float(redshift))
        self.device = device
        self.tf_cb = self.tf_cbnu = None          # the reference's result globals, as arrays after every call
        self.growth_to_z0 = None

    def onek_hmpc(self, kk):
        """T_cb at wavenumbers kk [h Mpc^-1]; T_cbnu is left in `self.tf_cbnu` (the reference's global of that name)."""
        k = _ffi.f64(np.atleast_1d(kk))
        cb, cbnu = np.empty(k.size), np.empty(k.size)
        g = C.c_double()
        a = self.args
        _ffi.check(_ffi.lib().lbg_eh_transfer(_ffi.context(self.device).handle, a[0], a[1], a[2], a[3], a[4], a[5], a[6], k.size,
                                              _ffi._dp(k.ravel()), _ffi._dp(cb), _ffi._dp(cbnu), C.byref(g)))
        self.tf_cb, self.tf_cbnu, self.growth_to_z0 = cb.reshape(k.shape), cbnu.reshape(k.shape), g.value
        return self.tf_cb if np.ndim(kk) else float(cb[0])

    def onek_mpc(self, kk):
        return self.onek_hmpc(np.asarray(kk, dtype=np.float64) / self.args[5])


def linear_power(k, omega_matter, omega_baryon, omega_lambda, hubble, ns, sigma8, omega_hdm=0.0, degen_hdm=1, device=None):
    """P_lin(k, z=0) [(Mpc/h)^3] on k [h/Mpc] = A k^ns T_cb^2 normalised to sigma8, built on the device; the table also stays on
    the context as the current linear table (what `lbg_upload_pk_linear` would have uploaded)."""
    k = _ffi.f64(k)
    w = _ffi.f64(simps_weights_x(np.log(k)))
    out = np.empty(k.size)
    ctx = _ffi.context(device)
    _ffi.check(_ffi.lib().lbg_eh_linear_table(ctx.handle, float(omega_matter), float(omega_baryon), float(omega_hdm), int(degen_hdm),
                                              float(omega_lambda), float(hubble), float(ns), float(sigma8), k.size, _ffi._dp(k),
                                              _ffi._dp(w), _ffi._dp(out)))
    _ffi.table_owner[ctx.device] = None          # nobody's python-side table: the next Cosmology call uploads its own
    return out
