// n0_modi.cu -- Hu-Okamoto N0 in the polar ("modi") form: replaces modi/nkk.py:35-92 (N_fisher: TT, EE, EB) and
// :94-146 (N_fisherTE).
//
//   int2(L) = sum_i w_l[i] l_i sum_j w_phi[j] frac(L, l_i, phi_j),   l2 = sqrt(L^2 + l^2 - 2 L l cos phi),
//   frac = 0 where l2 lies outside [l.min, l.max] (modi/nkk.py:84-85).
// Spectra are cubic splines on unit-spaced knots x0, x0+1, ... (CLASS multipoles); like scipy's
// InterpolatedUnivariateSpline(k=3) they extrapolate with the end polynomials, and the detector noise is
// part of the splined table (modi/nkk.py:209-210), not added analytically as in the root flavour.
//
// Work decomposition: thread <-> one l row (consecutive lanes = consecutive l, so the four B-spline rows a warp gathers at
// l2 lie within a few cache lines), loop over the phi nodes with cos/sin/weight broadcast from shared memory; the l-sum is
// a fixed-order block reduction followed by n0_modi_reduce_kernel (deterministic, no atomics).
#include "common.cuh"
#include <cmath>

namespace lbg {

struct ModiTab { const double* c; double x0; int n; };   // c[0..n+1]: uniform cubic B-spline coefficients c_{-1}..c_n

struct ModiParams {
    int n_L, n_l, n_phi, n_blk, mode;
    double lmin, lmax;
    ModiTab tab[4];
};

__device__ __forceinline__ double modi_spline(const ModiTab& T, double x) {
    const double s = x - T.x0;
    int i = (int)floor(s);
    i = max(0, min(i, T.n - 2));                       // end intervals extrapolate (ext=0 of FITPACK splev)
    const double t = s - (double)i, um = 1.0 - t;
    const double t2 = t * t, u2 = um * um;
    const double w0 = u2 * um * (1.0 / 6.0), w3 = t2 * t * (1.0 / 6.0);
    const double w1 = fma(0.5 * t2, t, (2.0 / 3.0) - t2), w2 = fma(0.5 * u2, um, (2.0 / 3.0) - u2);
    const double* c = T.c + i;
    return w0 * __ldg(c) + w1 * __ldg(c + 1) + w2 * __ldg(c + 2) + w3 * __ldg(c + 3);
}

constexpr int MODI_T = 128;

__global__ void __launch_bounds__(MODI_T)
n0_modi_kernel(ModiParams P, const double* __restrict__ Ls, const double* __restrict__ l, const double* __restrict__ w_l,
               const double* __restrict__ phi /*[3][n_phi]: cos, sin, weight*/, double* __restrict__ partial /*[n_L][n_blk]*/) {
    extern __shared__ double sh[];                     // [3][n_phi] then [MODI_T] reduction scratch
    double* red = sh + 3 * P.n_phi;
    for (int j = threadIdx.x; j < 3 * P.n_phi; j += MODI_T) sh[j] = phi[j];
    __syncthreads();
    const int r = blockIdx.x * MODI_T + threadIdx.x;
    const bool valid = r < P.n_l;
    const double L = Ls[blockIdx.y];
    double acc = 0.0;
    if (valid) {
        const double li = l[r];
        const double Ll = __dmul_rn(L, li);
        const double L2 = __dmul_rn(L, L), l2sq = __dmul_rn(li, li);
        const double base = __dadd_rn(L2, l2sq);
        const double u0 = modi_spline(P.tab[0], li), u1 = modi_spline(P.tab[1], li);
        const double u2 = modi_spline(P.tab[2], li), u3 = modi_spline(P.tab[3], li);
        for (int j = 0; j < P.n_phi; ++j) {
            const double c = sh[j], s = sh[P.n_phi + j], wj = sh[2 * P.n_phi + j];
            const double Llc = __dmul_rn(Ll, c);                                        // modi/nkk.py:50
            const double l2 = __dsqrt_rn(__dsub_rn(base, __dmul_rn(2.0, Llc)));        // :51
            if (l2 > P.lmax || l2 < P.lmin || !(l2 == l2)) continue;                    // :84-85 (NaN compares false there too,
                                                                                        //  but then poisons the sum; rounding can make
                                                                                        //  the radicand -0 only, never negative enough)
            double frac;
            if (P.mode == (int)LBG_TE) {                                                // modi/nkk.py:104-134
                const double Ll2c = L2 - Llc;
                const double s12 = L * s / l2, c12 = (L * c - li) / l2;
                const double c2 = c12 * c12 - s12 * s12;
                const double cl2te = modi_spline(P.tab[0], l2), cl2ntt = modi_spline(P.tab[1], l2);
                const double cl2nee = modi_spline(P.tab[2], l2), cl2nte = modi_spline(P.tab[3], l2);
                const double f12 = Llc * u0 * c2 + Ll2c * cl2te;
                const double f21 = Llc * u0 + Ll2c * cl2te * c2;
                const double xx = u3 * cl2nte;
                const double num = u2 * cl2ntt * f12 - xx * f21;
                const double den = u1 * cl2nee * u2 * cl2ntt - xx * xx;
                frac = f12 * (num / den);
            } else {                                                                    // modi/nkk.py:46-82
                const double cl2 = modi_spline(P.tab[1], l2), cln2 = modi_spline(P.tab[3], l2);
                const double a = Llc * u0 + (L2 - Llc) * cl2;
                double num = a * a, den = u2 * cln2;
                if (P.mode == (int)LBG_TT) {
                    den *= 2.0;
                } else {
                    const double s12 = L * s / l2, c12 = (L * c - li) / l2;
                    if (P.mode == (int)LBG_EB) {
                        const double t = 2.0 * s12 * c12;
                        num *= t * t;
                    } else {
                        const double t = c12 * c12 - s12 * s12;
                        num *= t * t;
                        den *= 2.0;
                    }
                }
                frac = num / den;
            }
            acc = fma(wj, frac, acc);
        }
        acc *= li * w_l[r];                                                             // :88-90
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int h = MODI_T / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[(size_t)blockIdx.y * P.n_blk + blockIdx.x] = red[0];
}

__global__ void __launch_bounds__(MODI_T)
n0_modi_reduce_kernel(int n_blk, const double* __restrict__ partial, double* __restrict__ out) {
    __shared__ double red[MODI_T];
    double acc = 0.0;
    for (int b = threadIdx.x; b < n_blk; b += MODI_T) acc += partial[(size_t)blockIdx.x * n_blk + b];
    red[threadIdx.x] = acc;
    __syncthreads();
    for (int h = MODI_T / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h) red[threadIdx.x] += red[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = red[0];
}

__global__ void modi_spline_eval_kernel(ModiTab T, int n, const double* __restrict__ x, double* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) out[t] = modi_spline(T, x[t]);
}
}  // namespace lbg

using namespace lbg;

// scipy InterpolatedUnivariateSpline(x0 + arange(n_knots), y, k=3)(x): extrapolates with the end polynomials (modi/nkk.py:205-213).
extern "C" int lbg_eval_spline_ext(lbg_ctx* c, int n_knots, double x0, const double* y, int n, const double* x, double* out) {
    LBG_REQUIRE(c && y && x && out, "lbg_eval_spline_ext: NULL argument");
    LBG_REQUIRE(n_knots >= 4, "lbg_eval_spline_ext: need at least 4 knots, got %d", n_knots);
    if (n <= 0) return LBG_OK;
    std::vector<double> coef((size_t)n_knots + 2);
    int rc = lbg_cubic_spline_coeffs(n_knots, y, coef.data());
    if (rc) return rc;
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    if (c->s0.reserve(coef.size() * 8) || c->s1.reserve((size_t)n * 8) || c->s2.reserve((size_t)n * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, coef.data(), coef.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, x, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    ModiTab T{c->s0.as<double>(), x0, n_knots};
    KernelTimer t(c);
    modi_spline_eval_kernel<<<(n + 255) / 256, 256, 0, c->stream>>>(T, n, c->s1.as<double>(), c->s2.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out, c->s2.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

extern "C" int lbg_n0_modi(lbg_ctx* c, int mode, int n_L, const double* L, int n_l, const double* l, const double* w_l,
                           int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                           const int* tab_n, const double* tab_x0, const double* tab_y, double* out_int) {
    LBG_REQUIRE(c && L && l && w_l && cos_phi && sin_phi && w_phi && tab_n && tab_x0 && tab_y && out_int, "lbg_n0_modi: NULL argument");
    LBG_REQUIRE(mode == (int)LBG_TT || mode == (int)LBG_EE || mode == (int)LBG_EB || mode == (int)LBG_TE, "lbg_n0_modi: mode %d is not one of TT, EE, EB, TE", mode);
    LBG_REQUIRE(n_L >= 0 && n_l > 0 && n_phi > 0 && n_phi <= 4096, "lbg_n0_modi: bad sizes n_L=%d n_l=%d n_phi=%d", n_L, n_l, n_phi);
    if (n_L == 0) return LBG_OK;
    size_t ncoef = 0;
    for (int t = 0; t < 4; ++t) {
        LBG_REQUIRE(tab_n[t] >= 4, "lbg_n0_modi: table %d has %d knots (need >= 4 for a not-a-knot cubic)", t, tab_n[t]);
        ncoef += (size_t)tab_n[t] + 2;
    }
    // not-a-knot cubic spline of each table in the uniform B-spline basis (same spline as FITPACK's k=3 interpolant)
    std::vector<double> coef(ncoef);
    {
        size_t oy = 0, oc = 0;
        for (int t = 0; t < 4; ++t) {
            int rc = lbg_cubic_spline_coeffs(tab_n[t], tab_y + oy, coef.data() + oc);
            if (rc) return rc;
            oy += tab_n[t]; oc += (size_t)tab_n[t] + 2;
        }
    }
    double lmin = l[0], lmax = l[0];
    for (int i = 1; i < n_l; ++i) { lmin = std::min(lmin, l[i]); lmax = std::max(lmax, l[i]); }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    ModiParams P;
    P.n_L = n_L; P.n_l = n_l; P.n_phi = n_phi; P.mode = mode; P.lmin = lmin; P.lmax = lmax;
    P.n_blk = (n_l + MODI_T - 1) / MODI_T;
    if (c->s0.reserve(ncoef * 8) || c->s1.reserve((size_t)n_L * 8) || c->s2.reserve((size_t)2 * n_l * 8) ||
        c->s3.reserve((size_t)3 * n_phi * 8) || c->s4.reserve((size_t)n_L * P.n_blk * 8) || c->s5.reserve((size_t)n_L * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    {
        size_t oc = 0;
        for (int t = 0; t < 4; ++t) {
            P.tab[t].c = c->s0.as<double>() + oc; P.tab[t].x0 = tab_x0[t]; P.tab[t].n = tab_n[t];
            oc += (size_t)tab_n[t] + 2;
        }
    }
    std::vector<double> ph(3 * (size_t)n_phi);
    memcpy(&ph[0], cos_phi, (size_t)n_phi * 8);
    memcpy(&ph[n_phi], sin_phi, (size_t)n_phi * 8);
    memcpy(&ph[2 * (size_t)n_phi], w_phi, (size_t)n_phi * 8);
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, coef.data(), ncoef * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, L, (size_t)n_L * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s2.p, l, (size_t)n_l * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s2.as<double>() + n_l, w_l, (size_t)n_l * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s3.p, ph.data(), ph.size() * 8, cudaMemcpyHostToDevice, st));
    const size_t smem = (3 * (size_t)n_phi + MODI_T) * 8;
    if (smem > 48 * 1024) LBG_CUDA(cudaFuncSetAttribute(n0_modi_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    KernelTimer t(c);
    n0_modi_kernel<<<dim3(P.n_blk, n_L), MODI_T, smem, st>>>(P, c->s1.as<double>(), c->s2.as<double>(), c->s2.as<double>() + n_l,
                                                              c->s3.as<double>(), c->s4.as<double>());
    t.launched();
    n0_modi_reduce_kernel<<<n_L, MODI_T, 0, st>>>(P.n_blk, c->s4.as<double>(), c->s5.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_int, c->s5.p, (size_t)n_L * 8, cudaMemcpyDeviceToHost, st));
    return t.finish();
}
