// ctx.cu -- context lifetime, error strings, CMB spline tables, FP64 peak micro-benchmark.
#include "common.cuh"
#include <future>
#include <algorithm>
#include <cstring>
#include <cstdarg>
#include <cmath>

namespace lbg {
static thread_local std::string g_err;
void set_error(const char* fmt, ...) {
    char buf[1024];
    va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
    g_err = buf;
}

// Not-a-knot cubic spline on unit-spaced knots 0..n-1, returned as the n+2 coefficients c[-1..n] of the
// uniform (cardinal) cubic B-spline basis:  S(x) = sum_j c_j B3(x - j).
// This is the spline scipy's interp1d(kind='cubic') = make_interp_spline(k=3) builds (prep_camb.py:88-89):
// a C2 piecewise cubic through the data whose third derivative is continuous at knots 1 and n-2.
// With M_i = S''(i):  M_{i-1} + 4 M_i + M_{i+1} = 6 d_i  (d_i = second difference of y), the not-a-knot
// conditions give M_1 = d_1 and M_{n-2} = d_{n-2};  c_i = y_i - M_i / 6.
void cardinal_bspline_coeffs(int n, const double* y, std::vector<double>& c /* n+2, c[j+1] = c_j */) {
    c.assign(n + 2, 0.0);
    std::vector<double> M(n, 0.0);
    if (n >= 4) {
        auto d = [&](int i) { return (y[i - 1] - 2.0 * y[i]) + y[i + 1]; };
        M[1] = d(1);
        M[n - 2] = d(n - 2);
        int m = n - 4;  // unknowns M_2 .. M_{n-3}
        if (m > 0) {
            std::vector<double> cp(m), dp(m);
            for (int k = 0; k < m; ++k) {
                int i = k + 2;
                double rhs = 6.0 * d(i);
                if (k == 0) rhs -= M[1];
                if (k == m - 1) rhs -= M[n - 2];
                double denom = 4.0 - (k ? cp[k - 1] : 0.0);
                cp[k] = 1.0 / denom;
                dp[k] = (rhs - (k ? dp[k - 1] : 0.0)) / denom;
            }
            for (int k = m - 1; k >= 0; --k) M[k + 2] = dp[k] - (k < m - 1 ? cp[k] * M[k + 3] : 0.0);
        }
        M[0] = 2.0 * M[1] - M[2];
        M[n - 1] = 2.0 * M[n - 2] - M[n - 3];
    }
    for (int i = 0; i < n; ++i) c[i + 1] = y[i] - M[i] / 6.0;
    if (n >= 2) {
        c[0] = 2.0 * c[1] - c[2] + M[0];
        c[n + 1] = 2.0 * c[n] - c[n - 1] + M[n - 1];
    }
}

// scipy's own representation of the same interpolant: make_interp_spline(arange(n), y, k=3) -> knots t = (0,0,0,0,2,3,..,n-3,
// n-1 x 4) (`_not_a_knot`), coefficients from the collocation matrix (`_coloc`, de Boor) solved by LAPACK dgbsv with kl = ku = 3.
// The sums below follow dgbtf2 + dgbtrs + dtbsv statement by statement, with the rank-1 / axpy updates fused (fma) and the
// multipliers formed by the reciprocal of the pivot, which is how the OpenBLAS kernels scipy ships compute them on FMA-capable
// x86-64: the coefficients come out bit-identical to scipy's (tests/test_scipy_spline_cpu.py).  Needed only where the reference's
// result hangs on the rounding of this routine (n0.cu, rows below the first non-zero C_l).
void scipy_notaknot_cubic(int n, const double* y, double* t_out, double* c_out) {
    const int k = 3, kl = 3, ku = 3, kv = kl + ku, ldab = 2 * kl + ku + 1;
    std::vector<double> t((size_t)n + k + 1);
    for (int i = 0; i < 4; ++i) t[i] = 0.0;
    for (int i = 2; i <= n - 3; ++i) t[4 + i - 2] = (double)i;
    for (int i = 0; i < 4; ++i) t[n + i] = (double)(n - 1);
    if (t_out) memcpy(t_out, t.data(), t.size() * 8);
    std::vector<double> AB((size_t)ldab * n, 0.0);
    auto ab = [&](int i, int j) -> double& { return AB[(size_t)j * ldab + i]; };
    int left = k;
    for (int j = 0; j < n; ++j) {                       // collocation rows: the four non-zero B-splines at x = j
        const double xv = (double)j;
        while (left < n - 1 && xv >= t[left + 1]) ++left;
        double h[4], hh[4];
        h[0] = 1.0;
        for (int q = 1; q <= k; ++q) {
            memcpy(hh, h, q * sizeof(double));
            h[0] = 0.0;
            for (int m = 1; m <= q; ++m) {
                const int ind = left + m;
                const double xb = t[ind], xa = t[ind - q];
                if (xb == xa) { h[m] = 0.0; continue; }
                const double w = hh[m - 1] / (xb - xa);
                h[m - 1] += w * (xb - xv);
                h[m] = w * (xv - xa);
            }
        }
        for (int a = 0; a <= k; ++a) { const int clmn = left - k + a; ab(kl + ku + j - clmn, clmn) = h[a]; }
    }
    std::vector<double> b(y, y + n);
    std::vector<int> ipiv(n);
    int ju = 0;
    for (int j = 0; j < n; ++j) {                       // dgbtf2
        const int km = std::min(kl, n - 1 - j);
        int jp = 0;
        double mx = fabs(ab(kv, j));
        for (int i = 1; i <= km; ++i) if (fabs(ab(kv + i, j)) > mx) { mx = fabs(ab(kv + i, j)); jp = i; }
        ipiv[j] = jp + j;
        if (ab(kv + jp, j) == 0.0) continue;            // singular column: cannot happen for distinct knots
        ju = std::max(ju, std::min(j + ku + jp, n - 1));
        if (jp != 0) for (int q = 0; q <= ju - j; ++q) std::swap(ab(kv + jp - q, j + q), ab(kv - q, j + q));
        if (km > 0) {
            const double r = 1.0 / ab(kv, j);
            for (int i = 1; i <= km; ++i) ab(kv + i, j) *= r;
            for (int q = 1; q <= ju - j; ++q) {
                const double yj = ab(kv - q, j + q);
                if (yj != 0.0) { const double temp = -yj; for (int i = 1; i <= km; ++i) ab(kv - q + i, j + q) = fma(ab(kv + i, j), temp, ab(kv - q + i, j + q)); }
            }
        }
    }
    for (int j = 0; j < n - 1; ++j) {                   // dgbtrs, L
        const int lm = std::min(kl, n - 1 - j);
        if (ipiv[j] != j) std::swap(b[ipiv[j]], b[j]);
        const double temp = -b[j];
        for (int i = 1; i <= lm; ++i) b[j + i] = fma(ab(kv + i, j), temp, b[j + i]);
    }
    for (int j = n - 1; j >= 0; --j) {                  // dtbsv, U
        if (b[j] != 0.0) {
            b[j] = b[j] / ab(kv, j);
            const double temp = b[j];
            for (int i = j - 1; i >= std::max(0, j - kv); --i) b[i] = fma(-temp, ab(kv + i - j, j), b[i]);
        }
    }
    memcpy(c_out, b.data(), (size_t)n * 8);
}

// BSpline.__call__ of that representation at one point (`_bspl.evaluate_spline` -> `find_interval`, `_deBoor_D`, k = 3, nu = 0), in its
// operation order; interp1d(bounds_error=False, fill_value=0.0) semantics outside [0, n-1] are the caller's.
double scipy_bspline_eval(const double* t, const double* c, int n, double x) {
    const int k = 3;
    int ell = k;
    while (x >= t[ell + 1] && ell + 1 != n) ++ell;
    double hh[4], h[4];
    h[0] = 1.0;
    for (int j = 1; j <= k; ++j) {
        memcpy(hh, h, j * sizeof(double));
        h[0] = 0.0;
        for (int q = 1; q <= j; ++q) {
            const int ind = ell + q;
            const double xb = t[ind], xa = t[ind - j];
            if (xb == xa) { h[q] = 0.0; continue; }
            const double w = hh[q - 1] / (xb - xa);
            h[q - 1] += w * (xb - x);
            h[q] = w * (x - xa);
        }
    }
    double out = 0.0;
    for (int a = 0; a <= k; ++a) out += c[ell + a - k] * h[a];
    return out;
}

int ensure_scipy_bspline(lbg_ctx* c) {
    if (c->cl_bspl_gen == c->cl_gen && !c->cl_bspl_t.empty()) return LBG_OK;
    const int n = c->n_ell;
    if (n < 6 || c->cl_host.size() != (size_t)8 * n) { set_error("scipy-order spline needs the host copy of at least 6 knots"); return LBG_ESTATE; }
    c->cl_bspl_t.assign((size_t)n + 4, 0.0);
    c->cl_bspl_c.assign((size_t)8 * n, 0.0);
    std::vector<std::future<void>> jobs;
    for (int s = 0; s < 8; ++s)
        jobs.push_back(std::async(std::launch::async, [c, n, s] {
            scipy_notaknot_cubic(n, c->cl_host.data() + (size_t)s * n, s == 0 ? c->cl_bspl_t.data() : nullptr, c->cl_bspl_c.data() + (size_t)s * n);
        }));
    for (auto& j : jobs) j.get();
    c->cl_bspl_gen = c->cl_gen;
    return LBG_OK;
}

__global__ void eval_cl_kernel(const ClRow* __restrict__ rows, int n_ell, int col, int n,
                               const double* __restrict__ x, double* __restrict__ out) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double xv = x[t], r = 0.0;
    if (xv >= 0.0 && xv <= (double)(n_ell - 1)) {
        int i = (int)xv;
        if (i > n_ell - 2) i = n_ell - 2;
        double tt = xv - (double)i, um = 1.0 - tt;
        double w0 = um * um * um * (1.0 / 6.0), w3 = tt * tt * tt * (1.0 / 6.0);
        double w1 = (2.0 / 3.0) - tt * tt + 0.5 * tt * tt * tt;
        double w2 = (2.0 / 3.0) - um * um + 0.5 * um * um * um;
        const double* p = reinterpret_cast<const double*>(rows + i) + col;
        r = w0 * p[0] + w1 * p[8] + w2 * p[16] + w3 * p[24];
    }
    out[t] = r;
}

// 8 independent DFMA chains per thread, 4096 iterations: FP64-pipe bound.
__global__ void dfma_peak_kernel(double* out, double a, double b, int iters) {
    double x0 = threadIdx.x, x1 = x0 + 1, x2 = x0 + 2, x3 = x0 + 3, x4 = x0 + 4, x5 = x0 + 5, x6 = x0 + 6, x7 = x0 + 7;
    for (int i = 0; i < iters; ++i) {
        x0 = fma(x0, a, b); x1 = fma(x1, a, b); x2 = fma(x2, a, b); x3 = fma(x3, a, b);
        x4 = fma(x4, a, b); x5 = fma(x5, a, b); x6 = fma(x6, a, b); x7 = fma(x7, a, b);
    }
    double s = ((x0 + x1) + (x2 + x3)) + ((x4 + x5) + (x6 + x7));
    if (s == 12345.678) out[0] = s;   // never true; keeps the chains alive
}
}  // namespace lbg

using namespace lbg;

extern "C" const char* lbg_last_error(void) { return g_err.c_str(); }
extern "C" const char* lbg_version(void) { return "lbgcmb-b200 0.1.0 (sm_100a)"; }

extern "C" int lbg_ctx_create(int device, lbg_ctx** out) { return lbg_ctx_create_lane(device, 0, out); }

extern "C" int lbg_ctx_create_lane(int device, int lane, lbg_ctx** out) {
    LBG_REQUIRE(out != nullptr, "lbg_ctx_create: out is NULL");
    LBG_REQUIRE(lane >= 0, "lbg_ctx_create_lane: lane %d is negative", lane);
    *out = nullptr;
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0) {
        set_error("lbg_ctx_create: no CUDA device (%s); this library has no CPU fallback",
                  e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        return LBG_ECUDA;
    }
    LBG_REQUIRE(device >= 0 && device < n, "lbg_ctx_create: device %d out of range [0,%d)", device, n);
    LBG_CUDA(cudaSetDevice(device));
    lbg_ctx* c = new lbg_ctx();
    c->device = device;
    c->lane = lane;
    LBG_CUDA(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, device));
    // lane 0 gets the device's highest stream priority, any other lane the lowest: when both lanes have CTAs pending, the short
    // lane-0 kernels (Limber, covariance) are dispatched first as SM slots free up instead of waiting for the whole N0 grid
    int prio_least = 0, prio_greatest = 0;
    LBG_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    LBG_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, lane == 0 ? prio_greatest : prio_least));
    LBG_CUDA(cudaEventCreate(&c->ev0));
    LBG_CUDA(cudaEventCreate(&c->ev1));
    LBG_CUDA(cudaEventCreate(&c->ev2));
    LBG_CUDA(cudaEventCreate(&c->ev3));
    *out = c;
    return LBG_OK;
}

void lbg_n0_release(lbg_ctx* c);   // n0.cu
void lbg_hf_release(lbg_ctx* c);   // halofit.cu
void lbg_xchg_release(lbg_ctx* c); // xchg.cu

extern "C" int lbg_ctx_destroy(lbg_ctx* c) {
    if (!c) return LBG_OK;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    lbg_n0_release(c);
    lbg_hf_release(c);
    lbg_xchg_release(c);
    for (DevBuf* b : {&c->cl_rows, &c->cl_soa, &c->cl_poly, &c->n0_L, &c->n0_utab, &c->n0_partial, &c->n0_out, &c->n0_plan_dev,
                      &c->s0, &c->s1, &c->s2, &c->s3, &c->s4, &c->s5, &c->pk_lnk, &c->pk_plin, &c->pk_pnl, &c->pk_rdk})
        b->release();
    for (DevBuf& b : c->pk_store) b.release();
    for (DevBuf& b : c->spec_set) b.release();
    c->pin_in.release(); c->pin_out.release(); c->pin_n0.release(); c->pin_n0_out.release();
    cudaEventDestroy(c->ev0); cudaEventDestroy(c->ev1); cudaEventDestroy(c->ev2); cudaEventDestroy(c->ev3);
    cudaStreamDestroy(c->stream);
    delete c;
    return LBG_OK;
}

extern "C" int lbg_last_kernel_ms(lbg_ctx* c, double* ms, int* launches) {
    LBG_REQUIRE(c, "lbg_last_kernel_ms: ctx is NULL");
    if (ms) *ms = c->last_ms;
    if (launches) *launches = c->last_launches;
    return LBG_OK;
}

extern "C" int lbg_last_main_kernel_ms(lbg_ctx* c, double* ms) {
    LBG_REQUIRE(c && ms, "lbg_last_main_kernel_ms: NULL argument");
    *ms = c->last_main_ms;
    return LBG_OK;
}

extern "C" int lbg_kernel_totals(lbg_ctx* c, double* ms, long long* launches, int reset) {
    LBG_REQUIRE(c, "lbg_kernel_totals: ctx is NULL");
    std::lock_guard<std::mutex> g(c->mu);
    if (ms) *ms = c->total_ms;
    if (launches) *launches = c->total_launches;
    if (reset) { c->total_ms = 0.0; c->total_launches = 0; }
    return LBG_OK;
}

extern "C" int lbg_fp64_peak(lbg_ctx* c, double* tflops) {
    LBG_REQUIRE(c && tflops, "lbg_fp64_peak: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    if (c->s0.reserve(64)) return LBG_ENOMEM;
    const int iters = 4096, threads = 256, blocks = c->sm_count * 16;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        KernelTimer t(c);
        dfma_peak_kernel<<<blocks, threads, 0, c->stream>>>(c->s0.as<double>(), 1.0000001, 1e-9, iters);
        t.launched();
        int rc = t.finish();
        if (rc) return rc;
        double flop = 2.0 * 8.0 * iters * (double)threads * blocks;
        double tf = flop / (c->last_ms * 1e-3) * 1e-12;
        if (rep > 0 && tf > best) best = tf;
    }
    *tflops = best;
    return LBG_OK;
}

extern "C" int lbg_upload_cl_tables(lbg_ctx* c, int n_ell, const double* lensed, const double* unlensed) {
    LBG_REQUIRE(c && lensed && unlensed, "lbg_upload_cl_tables: NULL argument");
    LBG_REQUIRE(n_ell >= 4, "lbg_upload_cl_tables: need at least 4 knots, got %d", n_ell);
    std::lock_guard<std::mutex> g(c->mu);
    if (c->n0_pending) { set_error("lbg_upload_cl_tables: an N0 call is pending on this context"); return LBG_ESTATE; }
    LBG_CUDA(cudaSetDevice(c->device));
    // input order TT,EE,BB,TE (prep_camb.py:84) -> row order TT,TE,EE,BB
    const int src_of_col[4] = {0, 3, 1, 2};
    // A forecast loop hands over the same tables call after call (only beam / noise / L change): compare with the resident copy,
    // value for value, and keep the splines when nothing changed (the whole input is still read once per call).
    if (c->n_ell == n_ell && c->cl_host.size() == (size_t)8 * n_ell) {
        bool same = true;
        for (int which = 0; which < 2 && same; ++which)
            for (int col = 0; col < 4 && same; ++col)
                same = !memcmp(c->cl_host.data() + (size_t)(which * 4 + col) * n_ell,
                               (which ? lensed : unlensed) + (size_t)src_of_col[col] * n_ell, (size_t)n_ell * 8);
        if (same) return LBG_OK;
    }
    std::vector<ClRow> rows(n_ell + 2);
    // the eight spline solves are independent sequential recurrences (0.1 ms each at 10^4 knots): one host thread per spectrum,
    // each writing its coefficients straight into the spectrum-major table the N0 main kernel reads (neighbouring lanes =
    // neighbouring |v| then read neighbouring doubles)
    const int stride = (n_ell + 2 + 15) & ~15;
    std::vector<double> soa((size_t)8 * stride, 0.0);
    {
        std::vector<std::future<void>> jobs;
        for (int which = 0; which < 2; ++which)
            for (int col = 0; col < 4; ++col)
                jobs.push_back(std::async(std::launch::async, [&, which, col] {
                    std::vector<double> coef;
                    cardinal_bspline_coeffs(n_ell, (which ? lensed : unlensed) + (size_t)src_of_col[col] * n_ell, coef);
                    memcpy(&soa[(size_t)(which * 4 + col) * stride], coef.data(), (size_t)(n_ell + 2) * 8);
                }));
        for (auto& j : jobs) j.get();
    }
    for (int r = 0; r < n_ell + 2; ++r)
        for (int k = 0; k < 8; ++k) reinterpret_cast<double*>(&rows[r])[k] = soa[(size_t)k * stride + r];
    if (c->cl_rows.reserve(rows.size() * sizeof(ClRow)) || c->cl_soa.reserve(soa.size() * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->cl_rows.p, rows.data(), rows.size() * sizeof(ClRow), cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->cl_soa.p, soa.data(), soa.size() * 8, cudaMemcpyHostToDevice, c->stream));
    c->cl_soa_stride = stride;
    LBG_CUDA(cudaStreamSynchronize(c->stream));
    c->n_ell = n_ell;
    c->ub_zero = true;
    for (int r = 0; r < n_ell; ++r) c->ub_zero = c->ub_zero && unlensed[(size_t)2 * n_ell + r] == 0.0;
    c->cl_host.resize((size_t)8 * n_ell);
    for (int which = 0; which < 2; ++which)
        for (int col = 0; col < 4; ++col)
            memcpy(c->cl_host.data() + (size_t)(which * 4 + col) * n_ell, (which ? lensed : unlensed) + (size_t)src_of_col[col] * n_ell, (size_t)n_ell * 8);
    ++c->cl_gen;
    return LBG_OK;
}

extern "C" int lbg_eval_cl(lbg_ctx* c, int which, int spec, int n, const double* x, double* out) {
    LBG_REQUIRE(c && x && out, "lbg_eval_cl: NULL argument");
    LBG_REQUIRE(c->n_ell > 0, "lbg_eval_cl: no C_l tables uploaded");
    LBG_REQUIRE((which == 0 || which == 1) && spec >= 0 && spec < 4, "lbg_eval_cl: bad which/spec %d/%d", which, spec);
    if (n <= 0) return LBG_OK;
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const int col_of_spec[4] = {0, 2, 3, 1};   // TT,EE,BB,TE -> TT,TE,EE,BB columns
    int col = (which == 0 ? 4 : 0) + col_of_spec[spec];
    if (c->s0.reserve((size_t)n * 8) || c->s1.reserve((size_t)n * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, x, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    KernelTimer t(c);
    eval_cl_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(c->cl_rows.as<ClRow>(), c->n_ell, col, n, c->s0.as<double>(), c->s1.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out, c->s1.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

// Stand-alone "interp1d(kind='cubic', bounds_error=False, fill_value=0.0)" on knots 0..n_knots-1.
extern "C" int lbg_eval_spline(lbg_ctx* c, int n_knots, const double* y, int n, const double* x, double* out) {
    LBG_REQUIRE(c && y && x && out, "lbg_eval_spline: NULL argument");
    LBG_REQUIRE(n_knots >= 4, "lbg_eval_spline: need at least 4 knots, got %d", n_knots);
    if (n <= 0) return LBG_OK;
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    std::vector<double> coef;
    cardinal_bspline_coeffs(n_knots, y, coef);
    std::vector<ClRow> rows(n_knots + 2);
    memset(rows.data(), 0, rows.size() * sizeof(ClRow));
    for (int r = 0; r < n_knots + 2; ++r) rows[r].uTT = coef[r];
    if (c->s2.reserve(rows.size() * sizeof(ClRow)) || c->s0.reserve((size_t)n * 8) || c->s1.reserve((size_t)n * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s2.p, rows.data(), rows.size() * sizeof(ClRow), cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, x, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    KernelTimer t(c);
    eval_cl_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(c->s2.as<ClRow>(), n_knots, 0, n, c->s0.as<double>(), c->s1.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out, c->s1.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

// Host-only (no GPU): scipy `interp1d(arange(n_knots), y, kind='cubic', bounds_error=False, fill_value=0.0)(x)` in scipy's own
// operation order (LAPACK-order not-a-knot solve + de Boor evaluation): bit-identical values, for the places where the reference's
// result hangs on them.
extern "C" int lbg_scipy_cubic_eval_host(int n_knots, const double* y, int n, const double* x, double* out) {
    LBG_REQUIRE(y && x && out && n_knots >= 6 && n >= 0, "lbg_scipy_cubic_eval_host: bad argument");
    std::vector<double> t((size_t)n_knots + 4), cf((size_t)n_knots);
    scipy_notaknot_cubic(n_knots, y, t.data(), cf.data());
    const double xmax = (double)(n_knots - 1);
    for (int i = 0; i < n; ++i) out[i] = (x[i] >= 0.0 && x[i] <= xmax) ? scipy_bspline_eval(t.data(), cf.data(), n_knots, x[i]) : 0.0;
    return LBG_OK;
}

// Host-only helper (no GPU needed): the n_knots+2 cardinal B-spline coefficients c[-1..n_knots] the kernels use.
extern "C" int lbg_cubic_spline_coeffs(int n_knots, const double* y, double* coef_out) {
    LBG_REQUIRE(y && coef_out && n_knots >= 4, "lbg_cubic_spline_coeffs: bad argument");
    std::vector<double> coef;
    cardinal_bspline_coeffs(n_knots, y, coef);
    memcpy(coef_out, coef.data(), coef.size() * sizeof(double));
    return LBG_OK;
}
