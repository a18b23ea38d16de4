"""CPU restatement of the Eisenstein & Hu (1999) MDM transfer function (TEST INFRASTRUCTURE).

set_cosm     EisensteinHu/EisensteinAndHu.c:101-217 (TFmdm_set_cosm; theta_cmb = 2.7255/2.7 as the reference ships it, :122)
onek_mpc     :221-264 (TFmdm_onek_mpc)      onek_hmpc   :268-280
The reference computes in float32 with double-precision libm calls in between; this restatement is float64 throughout and is
pinned to the compiled reference (oracle/_ref/libeh_ref.so, built by oracle/Makefile from the file where it lies) at the
float32 level: relative 2e-6 (tests/golden/eh.npz holds the reference's own outputs so the check also runs on the GPU box).
`ref_transfer` calls the compiled reference through ctypes."""
import ctypes as C
import os

import numpy as np

REF_SO = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref', 'libeh_ref.so')


def set_cosm(omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble, redshift):
    c = {}
    theta_cmb = 2.7255 / 2.7
    degen_hdm = max(int(degen_hdm), 1)
    if omega_baryon <= 0:
        omega_baryon = 1e-5
    if omega_hdm <= 0:
        omega_hdm = 1e-5
    omega_curv = 1.0 - omega_matter - omega_lambda
    omhh = omega_matter * hubble ** 2
    obhh = omega_baryon * hubble ** 2
    f_baryon = omega_baryon / omega_matter
    f_hdm = omega_hdm / omega_matter
    f_cdm = 1.0 - f_baryon - f_hdm
    f_cb = f_cdm + f_baryon
    f_bnu = f_baryon + f_hdm
    z_equality = 25000.0 * omhh / theta_cmb ** 4
    z_drag_b1 = 0.313 * omhh ** -0.419 * (1 + 0.607 * omhh ** 0.674)
    z_drag_b2 = 0.238 * omhh ** 0.223
    z_drag = 1291 * omhh ** 0.251 / (1.0 + 0.659 * omhh ** 0.828) * (1.0 + z_drag_b1 * obhh ** z_drag_b2)
    y_drag = z_equality / (1.0 + z_drag)
    sound_horizon_fit = 44.5 * np.log(9.83 / omhh) / np.sqrt(1.0 + 10.0 * obhh ** 0.75)
    p_c = 0.25 * (5.0 - np.sqrt(1 + 24.0 * f_cdm))
    p_cb = 0.25 * (5.0 - np.sqrt(1 + 24.0 * f_cb))
    zp1 = 1.0 + redshift
    omega_denom = omega_lambda + zp1 ** 2 * (omega_curv + omega_matter * zp1)
    omega_lambda_z = omega_lambda / omega_denom
    omega_matter_z = omega_matter * zp1 ** 2 * zp1 / omega_denom
    growth_k0 = z_equality / zp1 * 2.5 * omega_matter_z / (omega_matter_z ** (4.0 / 7.0) - omega_lambda_z
                                                         + (1.0 + omega_matter_z / 2.0) * (1.0 + omega_lambda_z / 70.0))
    g0 = z_equality * 2.5 * omega_matter / (omega_matter ** (4.0 / 7.0) - omega_lambda
                                            + (1.0 + omega_matter / 2.0) * (1.0 + omega_lambda / 70.0))
    alpha_nu = f_cdm / f_cb * (5.0 - 2. * (p_c + p_cb)) / (5. - 4. * p_cb) * (1 + y_drag) ** (p_cb - p_c) \
        * (1 + f_bnu * (-0.553 + 0.126 * f_bnu * f_bnu)) \
        / (1 - 0.193 * np.sqrt(f_hdm * degen_hdm) + 0.169 * f_hdm * degen_hdm ** 0.2) \
        * (1 + (p_c - p_cb) / 2 * (1 + 1 / (3. - 4. * p_c) / (7. - 4. * p_cb)) / (1 + y_drag))
    c.update(theta_cmb=theta_cmb, num_degen_hdm=float(degen_hdm), omhh=omhh, f_hdm=f_hdm, f_cb=f_cb, p_cb=p_cb,
             growth_k0=growth_k0, growth_to_z0=growth_k0 / g0, sound_horizon_fit=sound_horizon_fit,
             alpha_gamma=np.sqrt(alpha_nu), beta_c=1 / (1 - 0.949 * f_bnu), hhubble=hubble,
             k_equality=0.0746 * omhh / theta_cmb ** 2, z_equality=z_equality, z_drag=z_drag)
    return c


def onek_mpc(c, kk):
    """(tf_cb, tf_cbnu) at wavenumbers kk [Mpc^-1]."""
    kk = np.asarray(kk, dtype=np.float64)
    qq = kk / c['omhh'] * c['theta_cmb'] ** 2
    f_hdm, nd, p_cb, g0 = c['f_hdm'], c['num_degen_hdm'], c['p_cb'], c['growth_k0']
    y_fs = 17.2 * f_hdm * (1 + 0.488 * f_hdm ** (-7.0 / 6.0)) * (nd * qq / f_hdm) ** 2
    temp1 = g0 ** (1.0 - p_cb)
    temp2 = (g0 / (1 + y_fs)) ** 0.7
    growth_cb = (1.0 + temp2) ** (p_cb / 0.7) * temp1
    growth_cbnu = (c['f_cb'] ** (0.7 / p_cb) + temp2) ** (p_cb / 0.7) * temp1
    gamma_eff = c['omhh'] * (c['alpha_gamma'] + (1 - c['alpha_gamma']) / (1 + (kk * c['sound_horizon_fit'] * 0.43) ** 4))
    qq_eff = qq * c['omhh'] / gamma_eff
    tf_sup_L = np.log(2.71828 + 1.84 * c['beta_c'] * c['alpha_gamma'] * qq_eff)
    tf_sup_C = 14.4 + 325 / (1 + 60.5 * qq_eff ** 1.11)
    tf_sup = tf_sup_L / (tf_sup_L + tf_sup_C * qq_eff ** 2)
    qq_nu = 3.92 * qq * np.sqrt(nd / f_hdm)
    max_fs = 1 + 1.2 * f_hdm ** 0.64 * nd ** (0.3 + 0.6 * f_hdm) / (qq_nu ** -1.6 + qq_nu ** 0.8)
    tf_master = tf_sup * max_fs
    return tf_master * growth_cb / g0, tf_master * growth_cbnu / g0


def onek_hmpc(c, kk):
    return onek_mpc(c, np.asarray(kk, dtype=np.float64) * c['hhubble'])


def linear_power(k, cosmo_args, ns, sigma8):
    """P(k) [(Mpc/h)^3] on k [h/Mpc]: k^ns T_cb(k)^2 normalised to sigma8 with a top-hat of 8 Mpc/h, trapezoid in ln k over the
    given grid (no reference counterpart: the reference takes P_lin from camb; this is the table builder of SURVEY.md 8(f) 1)."""
    k = np.asarray(k, dtype=np.float64)
    t = onek_hmpc(set_cosm(*cosmo_args), k)[0]
    p = k ** ns * t * t
    x = 8.0 * k
    w = np.where(x < 1e-3, 1.0 - x * x / 10.0, 3.0 * (np.sin(x) - x * np.cos(x)) / np.where(x == 0, 1.0, x) ** 3)
    f = p * k ** 3 / (2 * np.pi ** 2) * w * w
    lnk = np.log(k)
    s2 = np.sum(0.5 * (f[1:] + f[:-1]) * (lnk[1:] - lnk[:-1]))
    return p * (sigma8 ** 2 / s2)


def ref_available():
    return os.path.isfile(REF_SO)


def ref_transfer(cosmo_args, k_hmpc):
    """(tf_cb, tf_cbnu) from the compiled reference (float32), one wavenumber at a time as its API works."""
    lib = C.CDLL(REF_SO)
    lib.TFmdm_set_cosm.argtypes = [C.c_float, C.c_float, C.c_float, C.c_int, C.c_float, C.c_float, C.c_float]
    lib.TFmdm_set_cosm.restype = C.c_int
    lib.TFmdm_onek_hmpc.argtypes = [C.c_float]
    lib.TFmdm_onek_hmpc.restype = C.c_float
    om, ob, on, deg, ol, h, z = cosmo_args
    lib.TFmdm_set_cosm(om, ob, on, int(deg), ol, h, z)
    cbnu = C.c_float.in_dll(lib, 'tf_cbnu')
    out = np.empty((2, len(k_hmpc)))
    for i, kk in enumerate(k_hmpc):
        out[0, i] = lib.TFmdm_onek_hmpc(float(kk))
        out[1, i] = cbnu.value
    return out[0], out[1]
