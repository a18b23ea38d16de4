// sweep.cu -- the S/N + (sigma8, b1) Fisher / interloper-bias stage for a whole GRID of survey designs in one launch
// (BASELINE.json configs[4]: beam x noise x n(z) configurations).
//
// Per (noise configuration n, galaxy configuration g) this is the body of snr.py:118-124 (cumulative S/N of C_kg with
// Ckg.py:83-87 variance) and of interloper.py:200-260 (per-L Gaussian covariance of D = (kk, kg, gg), its inverse, the
// Fisher matrix of (sigma8, b1) and the bias vector of the contaminated spectra), i.e. exactly what
// lbg_gaussian_cov + lbg_fisher_accum + lbg_snr compute for one design -- same arithmetic per multipole (shared device
// functions), same fixed-order sums (thread t takes l = t, t + RED_T, ...; then the block tree of block_sum).
// One CTA per design: 210 noise levels x 64 n(z) = 13 440 CTAs of 256 threads (90 CTAs per SM), 4951 multipoles each;
// spectra are read once from L2 (C_gg / C_kg rows are shared by all noise levels, N_kk rows by all n(z)).
#include "common.cuh"
#include "fisher_dev.cuh"
#include <cmath>

namespace lbg {

struct GridCfg { double ngg, xngg, fid_b1, lmax_fisher, lmax_snr, pad0, pad1, pad2; };

__global__ void __launch_bounds__(RED_T)
forecast_grid_kernel(int n_ell, int n_cfg, const double* __restrict__ ell, const double* __restrict__ nmodes,
                     const double* __restrict__ ckk, const double* __restrict__ cgg, const double* __restrict__ ckg,
                     const double* __restrict__ xcgg, const double* __restrict__ xckg, const GridCfg* __restrict__ cfg,
                     const double* __restrict__ nkk /*[n_noise][n_ell]*/, double fsky, double fid_sig8,
                     double* __restrict__ out /*[n_noise][n_cfg][8]*/) {
    __shared__ double sm[7 * RED_T];
    const int ig = blockIdx.x, in = blockIdx.y;
    const GridCfg G = cfg[ig];
    const double* gg = cgg + (size_t)ig * n_ell;
    const double* kg = ckg + (size_t)ig * n_ell;
    const double* xgg = xcgg ? xcgg + (size_t)ig * n_ell : nullptr;
    const double* xkg = xckg ? xckg + (size_t)ig * n_ell : nullptr;
    const double* nk = nkk + (size_t)in * n_ell;
    double acc[7] = {0, 0, 0, 0, 0, 0, 0};           // snr^2, F00, F01, F10, F11, b0, b1
    for (int l = threadIdx.x; l < n_ell; l += RED_T) {
        const double L = ell[l], w = nmodes[l];
        // total spectra (C + N) of the two fields g, k  (interloper.py:38-40: N_kg = 0)
        const double Tgg = __dadd_rn(gg[l], G.ngg), Tkg = kg[l], Tkk = __dadd_rn(ckk[l], nk[l]);
        if (L <= G.lmax_snr) {                        // snr.py:22-23 with Ckg.py:83-87
            const double num = __dadd_rn(__dmul_rn(Tkk, Tgg), __dmul_rn(kg[l], kg[l]));
            const double den = __dmul_rn(__dadd_rn(__dmul_rn(2.0, L), 1.0), fsky);
            const double r = kg[l] / sqrt(num / den);
            acc[0] += w * (r * r);
        }
        if (L < G.lmax_fisher) {                      // interloper.py:221
            // fields: 0 = g, 1 = k; spectra in the order (kk, kg, gg) of interloper.py:232
            const double T[4] = {Tgg, Tkg, Tkg, Tkk};
            const int sa[3] = {1, 1, 0}, sb[3] = {1, 0, 0};
            double A[9], B[9], D[6];
            for (int i = 0; i < 3; ++i)
                for (int j = 0; j < 3; ++j) A[i * 3 + j] = gaussian_cov_entry(T, 2, sa[i], sb[i], sa[j], sb[j], L, fsky);
            const double DL[3] = {ckk[l], kg[l], gg[l]};
            const double s8 = 2.0 / fid_sig8;         // dp_DL, interloper.py:65-74
            D[0] = s8 * DL[0]; D[1] = s8 * DL[1]; D[2] = s8 * DL[2];
            D[3] = 0.0; D[4] = DL[1] / G.fid_b1; D[5] = 2.0 * DL[2] / G.fid_b1;
            for (int e = 0; e < 6; ++e) B[e] = D[e];
            const bool bias = xgg != nullptr;
            if (bias) {                               // interloper.py:224-227,250
                B[6] = 0.0;
                B[7] = kg[l] - xkg[l];
                B[8] = (gg[l] + G.ngg) - (xgg[l] + G.xngg);
            }
            solve_spd<3, 3>(3, bias ? 3 : 2, A, B);
            for (int p = 0; p < 2; ++p) {
                for (int r = 0; r < 2; ++r) {
                    double s = 0.0;
                    for (int i = 0; i < 3; ++i) s += D[p * 3 + i] * B[r * 3 + i];
                    acc[1 + p * 2 + r] += w * s;
                }
                if (bias) {
                    double s = 0.0;
                    for (int i = 0; i < 3; ++i) s += D[p * 3 + i] * B[6 + i];
                    acc[5 + p] += w * s;
                }
            }
        }
    }
    block_sum<7>(acc, sm);
    if (threadIdx.x == 0) {
        double* o = out + ((size_t)in * n_cfg + ig) * 8;
        o[0] = sqrt(acc[0]);
        for (int e = 1; e < 7; ++e) o[e] = acc[e];
        o[7] = 0.0;
    }
}
}  // namespace lbg

using namespace lbg;

extern "C" int lbg_forecast_grid(lbg_ctx* c, int n_ell, const double* ell, const double* nmodes, const double* ckk,
                                 int n_cfg, const double* cgg, const double* ckg, const double* xcgg, const double* xckg,
                                 const double* cfg /*[n_cfg][8]*/, int n_noise, const double* nkk, double fsky, double fid_sig8,
                                 double* out /*[n_noise][n_cfg][8]*/) {
    LBG_REQUIRE(c && ell && nmodes && ckk && cgg && ckg && cfg && nkk && out, "lbg_forecast_grid: NULL argument");
    LBG_REQUIRE((xcgg == nullptr) == (xckg == nullptr), "lbg_forecast_grid: give both contaminated spectra or neither");
    LBG_REQUIRE(n_ell > 0 && n_cfg > 0 && n_noise > 0 && n_noise <= 65535, "lbg_forecast_grid: bad sizes n_ell=%d n_cfg=%d n_noise=%d", n_ell, n_cfg, n_noise);
    LBG_REQUIRE(fsky > 0 && fid_sig8 > 0, "lbg_forecast_grid: fsky and fid_sig8 must be positive");
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    // every input travels in ONE pinned staging block: [ell | nmodes | ckk | cgg | ckg | (xcgg | xckg) | nkk | cfg]
    const int n_arr = xcgg ? 4 : 2;
    const size_t ne = (size_t)n_ell, ng = (size_t)n_cfg * ne, nn = (size_t)n_noise * ne, no = (size_t)n_noise * n_cfg * 8;
    const size_t n_in = 3 * ne + n_arr * ng + nn + (size_t)n_cfg * 8;
    if (c->s0.reserve(n_in * 8) || c->pin_in.reserve(n_in * 8) || c->s4.reserve(no * 8) || c->pin_out.reserve(no * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    double* h = c->pin_in.as<double>();
    size_t o = 0;
    auto put = [&](const double* src, size_t n) { size_t at = o; memcpy(h + o, src, n * 8); o += n; return at; };
    const size_t o_ell = put(ell, ne), o_nm = put(nmodes, ne), o_kk = put(ckk, ne), o_gg = put(cgg, ng), o_kg = put(ckg, ng);
    const size_t o_xgg = xcgg ? put(xcgg, ng) : 0, o_xkg = xcgg ? put(xckg, ng) : 0;
    const size_t o_nkk = put(nkk, nn), o_cfg = put(cfg, (size_t)n_cfg * 8);
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, h, o * 8, cudaMemcpyHostToDevice, st));
    const double* d = c->s0.as<double>();
    KernelTimer t(c);
    forecast_grid_kernel<<<dim3(n_cfg, n_noise), RED_T, 0, st>>>(
        n_ell, n_cfg, d + o_ell, d + o_nm, d + o_kk, d + o_gg, d + o_kg, xcgg ? d + o_xgg : nullptr, xcgg ? d + o_xkg : nullptr,
        reinterpret_cast<const GridCfg*>(d + o_cfg), d + o_nkk, fsky, fid_sig8, c->s4.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, c->s4.p, no * 8, cudaMemcpyDeviceToHost, st));
    const int rc = t.finish();
    if (rc == LBG_OK) memcpy(out, c->pin_out.p, no * 8);
    return rc;
}
