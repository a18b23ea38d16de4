// eh.cu -- Eisenstein & Hu (1999) CDM + baryon + massive-neutrino transfer function and the linear P(k) table built from it.
//
// Replaces: EisensteinHu/EisensteinAndHu.c:101-217 (TFmdm_set_cosm: ~30 scalars, computed here on the host in double),
//           :221-264 (TFmdm_onek_mpc: the per-wavenumber fit, one thread per k), :268-280 (TFmdm_onek_hmpc).
// The reference is float32 with global state; this is float64 and stateless (the scalars travel as a kernel argument).
// lbg_eh_linear_table additionally normalises k^ns T_cb^2 to sigma8 (top-hat R = 8 Mpc/h, trapezoid in ln k over the
// given grid) and leaves the table on the context exactly as lbg_upload_pk_linear would: a cosmology sweep needs no host table
// (SURVEY.md 8(f) item 1).
#include "common.cuh"
#include "fisher_dev.cuh"
#include <cmath>

namespace lbg {

struct EhCosm {
    double theta_cmb, num_degen_hdm, omhh, f_hdm, f_cb, p_cb, growth_k0, sound_horizon_fit, alpha_gamma, beta_c, hhubble;
};

// EisensteinAndHu.c:101-217 in double; same clamps (omega_baryon, omega_hdm <= 0 -> 1e-5; degen_hdm < 1 -> 1).
static int eh_set_cosm(double om, double ob, double ohdm, int degen, double ol, double h, double z, EhCosm& c, double* growth_to_z0) {
    if (!(h > 0.0) || !(z > -1.0) || !(om > 0.0)) return -1;
    c.theta_cmb = 2.7255 / 2.7;                                   // :122
    if (degen < 1) degen = 1;
    c.num_degen_hdm = (double)degen;
    if (ob <= 0) ob = 1e-5;
    if (ohdm <= 0) ohdm = 1e-5;
    const double omega_curv = 1.0 - om - ol;
    c.omhh = om * h * h;
    const double obhh = ob * h * h;
    const double f_baryon = ob / om;
    c.f_hdm = ohdm / om;
    const double f_cdm = 1.0 - f_baryon - c.f_hdm;
    c.f_cb = f_cdm + f_baryon;
    const double f_bnu = f_baryon + c.f_hdm;
    const double t2 = c.theta_cmb * c.theta_cmb;
    const double z_equality = 25000.0 * c.omhh / (t2 * t2);
    const double z_drag_b1 = 0.313 * pow(c.omhh, -0.419) * (1 + 0.607 * pow(c.omhh, 0.674));
    const double z_drag_b2 = 0.238 * pow(c.omhh, 0.223);
    const double z_drag = 1291 * pow(c.omhh, 0.251) / (1.0 + 0.659 * pow(c.omhh, 0.828)) * (1.0 + z_drag_b1 * pow(obhh, z_drag_b2));
    const double y_drag = z_equality / (1.0 + z_drag);
    c.sound_horizon_fit = 44.5 * log(9.83 / c.omhh) / sqrt(1.0 + 10.0 * pow(obhh, 0.75));
    const double p_c = 0.25 * (5.0 - sqrt(1 + 24.0 * f_cdm));
    c.p_cb = 0.25 * (5.0 - sqrt(1 + 24.0 * c.f_cb));
    const double zp1 = 1.0 + z;
    const double omega_denom = ol + zp1 * zp1 * (omega_curv + om * zp1);
    const double olz = ol / omega_denom, omz = om * zp1 * zp1 * zp1 / omega_denom;
    c.growth_k0 = z_equality / zp1 * 2.5 * omz / (pow(omz, 4.0 / 7.0) - olz + (1.0 + omz / 2.0) * (1.0 + olz / 70.0));
    const double g0 = z_equality * 2.5 * om / (pow(om, 4.0 / 7.0) - ol + (1.0 + om / 2.0) * (1.0 + ol / 70.0));
    if (growth_to_z0) *growth_to_z0 = c.growth_k0 / g0;
    const double alpha_nu = f_cdm / c.f_cb * (5.0 - 2. * (p_c + c.p_cb)) / (5. - 4. * c.p_cb) * pow(1 + y_drag, c.p_cb - p_c) *
                            (1 + f_bnu * (-0.553 + 0.126 * f_bnu * f_bnu)) /
                            (1 - 0.193 * sqrt(c.f_hdm * c.num_degen_hdm) + 0.169 * c.f_hdm * pow(c.num_degen_hdm, 0.2)) *
                            (1 + (p_c - c.p_cb) / 2 * (1 + 1 / (3. - 4. * p_c) / (7. - 4. * c.p_cb)) / (1 + y_drag));
    c.alpha_gamma = sqrt(alpha_nu);
    c.beta_c = 1 / (1 - 0.949 * f_bnu);
    c.hhubble = h;
    return 0;
}

// EisensteinAndHu.c:221-264 for wavenumber kk [Mpc^-1].
__device__ __forceinline__ void eh_onek_mpc(const EhCosm& c, double kk, double& tf_cb, double& tf_cbnu) {
    const double qq = kk / c.omhh * (c.theta_cmb * c.theta_cmb);
    const double r = c.num_degen_hdm * qq / c.f_hdm;
    const double y_fs = 17.2 * c.f_hdm * (1 + 0.488 * pow(c.f_hdm, -7.0 / 6.0)) * (r * r);
    const double temp1 = pow(c.growth_k0, 1.0 - c.p_cb);
    const double temp2 = pow(c.growth_k0 / (1 + y_fs), 0.7);
    const double growth_cb = pow(1.0 + temp2, c.p_cb / 0.7) * temp1;
    const double growth_cbnu = pow(pow(c.f_cb, 0.7 / c.p_cb) + temp2, c.p_cb / 0.7) * temp1;
    const double ks = kk * c.sound_horizon_fit * 0.43;
    const double ks2 = ks * ks;
    const double gamma_eff = c.omhh * (c.alpha_gamma + (1 - c.alpha_gamma) / (1 + ks2 * ks2));
    const double qq_eff = qq * c.omhh / gamma_eff;
    const double tf_sup_L = log(2.71828 + 1.84 * c.beta_c * c.alpha_gamma * qq_eff);
    const double tf_sup_C = 14.4 + 325 / (1 + 60.5 * pow(qq_eff, 1.11));
    const double tf_sup = tf_sup_L / (tf_sup_L + tf_sup_C * (qq_eff * qq_eff));
    const double qq_nu = 3.92 * qq * sqrt(c.num_degen_hdm / c.f_hdm);
    const double max_fs = 1 + 1.2 * pow(c.f_hdm, 0.64) * pow(c.num_degen_hdm, 0.3 + 0.6 * c.f_hdm) / (pow(qq_nu, -1.6) + pow(qq_nu, 0.8));
    const double tf_master = tf_sup * max_fs;
    tf_cb = tf_master * growth_cb / c.growth_k0;
    tf_cbnu = tf_master * growth_cbnu / c.growth_k0;
}

__global__ void eh_transfer_kernel(EhCosm c, int n, const double* __restrict__ k_hmpc, double* __restrict__ tcb, double* __restrict__ tcbnu) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double a, b;
    eh_onek_mpc(c, k_hmpc[i] * c.hhubble, a, b);                 // TFmdm_onek_hmpc, :268-280
    tcb[i] = a;
    tcbnu[i] = b;
}

// shape[i] = k^ns T_cb^2 ; integrand[i] = shape k^3 / (2 pi^2) W_TH(8k)^2
__global__ void eh_shape_kernel(EhCosm c, int n, double ns, const double* __restrict__ k, double* __restrict__ shape, double* __restrict__ f) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double kk = k[i];
    double a, b;
    eh_onek_mpc(c, kk * c.hhubble, a, b);
    const double p = pow(kk, ns) * a * a;
    const double x = 8.0 * kk;
    double w;
    if (x < 1e-3) w = 1.0 - x * x / 10.0;
    else { double s, co; sincos(x, &s, &co); w = 3.0 * (s - x * co) / (x * x * x); }
    shape[i] = p;
    f[i] = p * (kk * kk * kk) / (2.0 * 3.14159265358979323846 * 3.14159265358979323846) * w * w;
}

// sigma^2 = trapezoid of f over ln k (one block, fixed order); then P = shape * sigma8^2 / sigma^2 and the Delta^2 * weight row
__global__ void __launch_bounds__(RED_T)
eh_normalise_kernel(int n, double sigma8, const double* __restrict__ k, const double* __restrict__ shape, const double* __restrict__ f,
                    const double* __restrict__ w_lnk, double* __restrict__ plin, double* __restrict__ d2w) {
    __shared__ double sm[RED_T];
    double acc[1] = {0.0};
    for (int i = threadIdx.x; i + 1 < n; i += RED_T) acc[0] += 0.5 * (f[i + 1] + f[i]) * (log(k[i + 1]) - log(k[i]));
    block_sum<1>(acc, sm);
    __shared__ double scale;
    if (threadIdx.x == 0) scale = sigma8 * sigma8 / acc[0];
    __syncthreads();
    const double twopi2 = 2.0 * 3.14159265358979323846 * 3.14159265358979323846;
    for (int i = threadIdx.x; i < n; i += RED_T) {
        const double p = shape[i] * scale;
        plin[i] = p;
        d2w[i] = p * (k[i] * k[i] * k[i]) / twopi2 * w_lnk[i];
    }
}
}  // namespace lbg

using namespace lbg;

extern "C" int lbg_eh_transfer(lbg_ctx* c, double omega_matter, double omega_baryon, double omega_hdm, int degen_hdm, double omega_lambda,
                               double hubble, double redshift, int n_k, const double* k_hmpc, double* out_tf_cb, double* out_tf_cbnu,
                               double* out_growth_to_z0) {
    LBG_REQUIRE(c && k_hmpc && out_tf_cb && n_k > 0, "lbg_eh_transfer: bad argument");
    EhCosm ec;
    double g = 0.0;
    LBG_REQUIRE(eh_set_cosm(omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble, redshift, ec, &g) == 0,
                "lbg_eh_transfer: need hubble > 0, redshift > -1, omega_matter > 0 (EisensteinAndHu.c:135-148)");
    if (out_growth_to_z0) *out_growth_to_z0 = g;
    std::lock_guard<std::mutex> lk(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    if (c->s0.reserve((size_t)n_k * 8) || c->s1.reserve((size_t)2 * n_k * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, k_hmpc, (size_t)n_k * 8, cudaMemcpyHostToDevice, c->stream));
    double* o = c->s1.as<double>();
    KernelTimer t(c);
    eh_transfer_kernel<<<(n_k + 127) / 128, 128, 0, c->stream>>>(ec, n_k, c->s0.as<double>(), o, o + n_k);
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_tf_cb, o, (size_t)n_k * 8, cudaMemcpyDeviceToHost, c->stream));
    if (out_tf_cbnu) LBG_CUDA(cudaMemcpyAsync(out_tf_cbnu, o + n_k, (size_t)n_k * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

extern "C" int lbg_eh_linear_table(lbg_ctx* c, double omega_matter, double omega_baryon, double omega_hdm, int degen_hdm, double omega_lambda,
                                   double hubble, double ns, double sigma8, int n_k, const double* k, const double* w_lnk,
                                   double* out_plin /*or NULL*/) {
    LBG_REQUIRE(c && k && w_lnk, "lbg_eh_linear_table: NULL argument");
    LBG_REQUIRE(n_k >= 4 && n_k <= 6000, "lbg_eh_linear_table: n_k=%d outside [4, 6000]", n_k);
    LBG_REQUIRE(sigma8 > 0, "lbg_eh_linear_table: sigma8 must be positive");
    for (int i = 1; i < n_k; ++i) LBG_REQUIRE(k[i] > k[i - 1] && k[0] > 0, "lbg_eh_linear_table: k must be positive and increasing");
    EhCosm ec;
    LBG_REQUIRE(eh_set_cosm(omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble, 0.0, ec, nullptr) == 0,
                "lbg_eh_linear_table: need hubble > 0, omega_matter > 0");
    std::lock_guard<std::mutex> lk(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t nb = (size_t)n_k * 8;
    if (c->pk_lnk.reserve(3 * nb) || c->s0.reserve(nb) || c->s1.reserve(2 * nb)) return LBG_ENOMEM;
    double* tab = c->pk_lnk.as<double>();                        // [k | P_lin(z=0) | Delta^2 * weight], as lbg_upload_pk_linear leaves it
    LBG_CUDA(cudaMemcpyAsync(tab, k, nb, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, w_lnk, nb, cudaMemcpyHostToDevice, c->stream));
    double* sh = c->s1.as<double>();
    KernelTimer t(c);
    eh_shape_kernel<<<(n_k + 127) / 128, 128, 0, c->stream>>>(ec, n_k, ns, tab, sh, sh + n_k);
    t.launched();
    eh_normalise_kernel<<<1, RED_T, 0, c->stream>>>(n_k, sigma8, tab, sh, sh + n_k, c->s0.as<double>(), tab + n_k, tab + 2 * (size_t)n_k);
    t.launched();
    if (out_plin) LBG_CUDA(cudaMemcpyAsync(out_plin, tab + n_k, nb, cudaMemcpyDeviceToHost, c->stream));
    int rc = t.finish();
    if (rc) return rc;
    if ((rc = upload_rdk(c, n_k, k))) return rc;
    c->n_k = n_k;
    c->pk_npan0 = c->pk_npan1 = 0;           // no numpy-order Simpson data for a device-built table: sigma^2 by the weight-vector kernel
    c->drop_pk_tables();
    return LBG_OK;
}
