// limber.cu -- matter power spectrum table (linear table x growth -> HALOFIT) and Limber projections.
//
// Replaces: modi/halofit.py:36-44 (sigma^2(R) Simpson integrals), :139-212 (Takahashi/Smith formula),
//           modi/cosmology.py:158-182 (pkalin / pkanlin), and the Limber sums of Cgg.py:53-68, Ckg.py:42-59,
//           cmb/lensing.py:27-43, modi/lensing.py:127-179 (all "sum_z weight_z * P(k=(l+1/2)/chi_z, z)").
// The quintic-spline glue of modi/halofit.py:70-98 (k_nl, n_eff, curvature from 500 ln sigma samples) stays on the
// host with the same scipy routine the reference calls (SURVEY.md "hard parts"); everything per (z,k) or (l,z) is here.
#include "common.cuh"
#include "pk_kernels.cuh"
#include <cmath>
#include <algorithm>

namespace lbg {

// np.interp(x, k, P) with numpy's operation order; oob: 0 -> 0 outside [k0, k_{n-1}] (left=right=0), 1 -> clamp.
__device__ __forceinline__ double interp_k(const double* __restrict__ ks, const double* __restrict__ P, int n_k, double x, int oob) {
    if (x < ks[0]) return oob ? P[0] : 0.0;
    if (x > ks[n_k - 1]) return oob ? P[n_k - 1] : 0.0;
    if (x == ks[n_k - 1]) return P[n_k - 1];
    int lo = 0, hi = n_k - 1;                    // ks[lo] <= x < ks[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (ks[mid] <= x) lo = mid; else hi = mid;
    }
    const double slope = __ddiv_rn(__dsub_rn(P[lo + 1], P[lo]), __dsub_rn(ks[lo + 1], ks[lo]));
    return __dadd_rn(__dmul_rn(slope, __dsub_rn(x, ks[lo])), P[lo]);
}

#define LIMBER_BT 8
#define LIMBER_ZMIN 64     /* at least this many z nodes per split */
#define LIMBER_SMAX 30     /* 39 l-blocks (4951 multipoles) x 30 splits = 1170 CTAs: one wave of 148 SMs x 8 CTAs */
#ifndef LIMBER_U
#define LIMBER_U 4         /* z nodes in flight per thread: their table loads (L2 latency) overlap */
#endif
#define LIMBER_WALK 6      /* cells the bracket may move from the previous z node before the search falls back to bisection */
// Number of z splits: a function of n_z ONLY, so that a multipole's value does not depend on which other multipoles or
// batch entries travel in the same call (tests/test_limber_gpu.py::test_limber_full_grid_properties).
static inline int limber_splits(int n_z) { return std::max(1, std::min(LIMBER_SMAX, n_z / LIMBER_ZMIN)); }

// RN(a / b) from rb = RN(1 / b) (a true IEEE reciprocal, computed once per z node / per k cell on the host) in five FMA-pipe
// instructions instead of the ~45 of the division routine: q0 = RN(a rb) is within 1.5 ulp, the first residual step makes it
// faithful, and for a faithful q and a correctly rounded reciprocal the second step returns the correctly rounded quotient
// (Markstein 1990).  Outside the range where no intermediate can underflow / overflow it falls back to the division itself.
// Checked bit for bit against a / b on 4e9 random pairs and on the bench grid by scripts/micro/div_markstein.cu.
__device__ __forceinline__ double div_rn(double a, double b, double rb) {
    double q = a * rb;
    if (!(fabs(q) > 1e-250 && fabs(q) < 1e250)) return __ddiv_rn(a, b);
    double r = fma(-q, b, a);
    q = fma(r, rb, q);
    r = fma(-q, b, a);
    return fma(r, rb, q);
}

// The cell of np.interp: the unique lo with ks[lo] <= x < ks[lo + 1] (caller guarantees ks[0] <= x < ks[n_k - 1]).  x = (l + 1/2) / chi
// moves by a fraction of a cell between neighbouring z nodes, so the search starts from the previous node's cell and walks; same
// cell as the bisection of interp_k, hence the same value bit for bit.
__device__ __forceinline__ int hunt_cell(const double* __restrict__ ks, int n_k, double x, int lo) {
    int moved = 0;
    while (x < __ldg(ks + lo)) { --lo; if (++moved > LIMBER_WALK) break; }
    while (moved <= LIMBER_WALK && x >= __ldg(ks + lo + 1)) { ++lo; ++moved; }
    if (moved > LIMBER_WALK) {
        int hi = n_k - 1;
        lo = 0;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (__ldg(ks + mid) <= x) lo = mid; else hi = mid;
        }
    }
    return lo;
}

// partial[s][b][l] = sum_{z in split s} weight[b][z] * P_z(k = (l + 1/2) / chi_z).
// Block = 128 consecutive l x NB batch entries (they share each P lookup) x one z split.  The k grid and the reciprocal cell
// widths (32 KB together) are read through L1; the P(k,z) table rows stream from L2.  No shared memory, so the CTAs co-reside
// with whatever the other lane has on the SMs.
// Each trip handles LIMBER_U z nodes: cells first (short walks), then all table loads, then the interpolations and the sums
// IN z ORDER inside a split; split order in limber_reduce_kernel (deterministic, and the same additions in the same order as
// one node per trip).  NB is the number of batch entries of this launch's blocks (1 for the reference-style single calls, else
// LIMBER_BT with the last block's missing entries pointed at entry 0 and not stored).
template <int NB>
__global__ void __launch_bounds__(128, 8)
limber_kernel(int n_batch, int n_ell, int n_z, int n_k, int z_per_split, const double* __restrict__ ell, const double* __restrict__ zc /*chi | RN(1/chi)*/,
              const double* __restrict__ weight /*[n_batch][n_z]*/, const unsigned char* __restrict__ live_z /*[batch block][n_z]*/,
              const int* __restrict__ prow /*[n_z] table row per z node*/, const double* __restrict__ ks,
              const double* __restrict__ rdk /*[n_k - 1] RN(1 / (ks[i+1] - ks[i]))*/, const double* __restrict__ ptab, int oob,
              double* __restrict__ partial) {
    const int il = blockIdx.x * blockDim.x + threadIdx.x;
    if (il >= n_ell) return;
    const int b0 = blockIdx.y * NB;
    const int z0 = blockIdx.z * z_per_split, z1 = min(n_z, z0 + z_per_split);
    const double* __restrict__ chi = zc;
    const double* __restrict__ rchi = zc + n_z;
    const unsigned char* __restrict__ lv = live_z + (size_t)blockIdx.y * n_z;
    const double* wq[NB];
#pragma unroll
    for (int q = 0; q < NB; ++q) wq[q] = weight + (size_t)(b0 + q < n_batch ? b0 + q : 0) * n_z;
    const double lp = ell[il] + 0.5;
    const double k_first = __ldg(ks), k_last = __ldg(ks + n_k - 1);
    double acc[NB];
#pragma unroll
    for (int q = 0; q < NB; ++q) acc[q] = 0.0;
    int cell = (n_k - 1) >> 1;
    for (int iz = z0; iz < z1; iz += LIMBER_U) {
        bool live[LIMBER_U], inside[LIMBER_U];
        int lo[LIMBER_U];
        double x[LIMBER_U];
#pragma unroll
        for (int u = 0; u < LIMBER_U; ++u) {
            const int z = min(iz + u, z1 - 1);
            live[u] = (iz + u < z1) && lv[z];        // block-uniform
            x[u] = div_rn(lp, __ldg(chi + z), __ldg(rchi + z));   // Cgg.py:57, Ckg.py:51, cmb/lensing.py:34, modi/cosmology.py:156
            inside[u] = x[u] >= k_first && x[u] < k_last;
            lo[u] = x[u] < k_first ? 0 : n_k - 2;    // outside the grid, or exactly on the last knot: P[0] / P[n_k - 1] below
            if (live[u] && inside[u]) { cell = hunt_cell(ks, n_k, x[u], cell); lo[u] = cell; }
        }
        double p0[LIMBER_U], p1[LIMBER_U];
#pragma unroll
        for (int u = 0; u < LIMBER_U; ++u) {
            const double* row = ptab + (size_t)__ldg(prow + min(iz + u, z1 - 1)) * n_k + lo[u];
            p0[u] = live[u] ? __ldg(row) : 0.0;
            p1[u] = live[u] ? __ldg(row + 1) : 0.0;
        }
#pragma unroll
        for (int u = 0; u < LIMBER_U; ++u) {
            if (!live[u]) continue;
            double pv;
            if (inside[u]) {                         // np.interp with numpy's operation order
                const double k0 = __ldg(ks + lo[u]), k1 = __ldg(ks + lo[u] + 1);
                const double slope = div_rn(__dsub_rn(p1[u], p0[u]), __dsub_rn(k1, k0), __ldg(rdk + lo[u]));
                pv = __dadd_rn(__dmul_rn(slope, __dsub_rn(x[u], k0)), p0[u]);
            } else if (x[u] < k_first) pv = oob ? p0[u] : 0.0;
            else if (x[u] > k_last) pv = oob ? p1[u] : 0.0;
            else pv = x[u] == k_last ? p1[u] : x[u] + p1[u];   // exactly the last knot; a NaN node (chi = NaN) stays NaN as in np.interp
            const int z = iz + u;
#pragma unroll
            for (int q = 0; q < NB; ++q) acc[q] += __ldg(wq[q] + z) * pv;
        }
    }
#pragma unroll
    for (int q = 0; q < NB; ++q)
        if (b0 + q < n_batch) partial[((size_t)blockIdx.z * n_batch + b0 + q) * n_ell + il] = acc[q];
}

__global__ void limber_reduce_kernel(size_t n, int n_split, const double* __restrict__ partial, double* __restrict__ out) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double v = partial[t];
    for (int s = 1; s < n_split; ++s) v += partial[(size_t)s * n + t];
    out[t] = v;
}

__global__ void pk_eval_kernel(int n, int n_k, const double* __restrict__ ks, const double* __restrict__ prow_ptr,
                               const double* __restrict__ x, int oob, double* __restrict__ out) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) out[t] = interp_k(ks, prow_ptr, n_k, x[t], oob);
}

// ---- fused forecast spectra (lbg_forecast_spectra) ---------------------------------------------------------------------------
// Per-node Limber weights of one galaxy sample on one z grid, built on the device from the sampled p(z), b(z):
//   sliced_pz.py:4-27   p^ = p / norm, zero outside [zmin, zmax], then divided by sum(p^) * dz
//   kind 0 (C_gg)       Cgg.py:53-68   w = simps_w * (H/c) p^ p^ / chi^2 * b b / h
//   kind 1 (C_kg)       Ckg.py:42-57   w = simps_w * W_kappa p^ / chi^2 * b / h^2
// One block per (sample, kind); the sum over the grid is a fixed-order block reduction.
struct WeightGrid { int n; const double *z, *chi, *fac, *w, *p, *b; double* out; double dz; };

__global__ void __launch_bounds__(256)
forecast_weights_kernel(WeightGrid gg, WeightGrid kg, const double* __restrict__ norm, double zmin, double zmax, double h100) {
    __shared__ double sm[256];
    const int s = blockIdx.x, kind = blockIdx.y;
    const WeightGrid G = kind == 0 ? gg : kg;
    const double* p = G.p + (size_t)s * G.n;
    const double* b = G.b + (size_t)s * G.n;
    const double nrm = norm[s];
    double acc = 0.0;
    for (int i = threadIdx.x; i < G.n; i += 256) {
        const double z = G.z[i];
        double v = p[i] / nrm;
        if (z < zmin || z > zmax) v = 0.0;
        acc += v;
    }
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int hh = 128; hh > 0; hh >>= 1) {
        if (threadIdx.x < hh) sm[threadIdx.x] += sm[threadIdx.x + hh];
        __syncthreads();
    }
    const double den = sm[0] * G.dz;
    for (int i = threadIdx.x; i < G.n; i += 256) {
        const double z = G.z[i], chi = G.chi[i];
        double v = p[i] / nrm;
        if (z < zmin || z > zmax) v = 0.0;
        v = v / den;
        double wv;
        if (kind == 0) wv = (((G.w[i] * (((G.fac[i] * v) * v) / (chi * chi))) * b[i]) * b[i]) / h100;
        else           wv = ((G.w[i] * ((G.fac[i] * v) / (chi * chi))) * b[i]) / (h100 * h100);
        G.out[(size_t)s * G.n + i] = wv;
    }
}

// live[bb][z] = some batch entry of batch block bb weights z node z (limber_kernel skips the others)
__global__ void limber_live_kernel(int n_batch, int nbk, int n_z, const double* __restrict__ weight, unsigned char* __restrict__ live) {
    const int z = blockIdx.x * blockDim.x + threadIdx.x, bb = blockIdx.y;
    if (z >= n_z) return;
    bool any = false;
    for (int q = bb * nbk; q < min(n_batch, (bb + 1) * nbk); ++q) any = any || (weight[(size_t)q * n_z + z] != 0.0);
    live[(size_t)bb * n_z + z] = any;
}

__global__ void iota_kernel(int n, int* out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = i;
}
}  // namespace lbg

using namespace lbg;

// RN(1 / (k[i+1] - k[i])) for the Limber kernel's np.interp slope (div_rn); called by everything that installs a k grid.
int lbg::upload_rdk(lbg_ctx* c, int n_k, const double* k) {
    std::vector<double> r((size_t)n_k, 0.0);
    for (int i = 0; i + 1 < n_k; ++i) r[i] = 1.0 / (k[i + 1] - k[i]);
    if (c->pk_rdk.reserve(r.size() * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->pk_rdk.p, r.data(), r.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaStreamSynchronize(c->stream));
    return LBG_OK;
}

// lnk (may be NULL: log(k) of this libm): the ln k values the caller's numpy computed, so that the numpy-order sigma^2 sums
// (sigma2_np_kernel) see the reference's own panel widths.
extern "C" int lbg_upload_pk_linear2(lbg_ctx* c, int n_k, const double* k, const double* plin_z0, const double* w_lnk, const double* lnk) {
    LBG_REQUIRE(c && k && plin_z0 && w_lnk, "lbg_upload_pk_linear: NULL argument");
    LBG_REQUIRE(n_k >= 4 && n_k <= 6000, "lbg_upload_pk_linear: n_k=%d outside [4, 6000] (k grid is staged in shared memory)", n_k);
    for (int i = 1; i < n_k; ++i) LBG_REQUIRE(k[i] > k[i - 1] && k[0] > 0, "lbg_upload_pk_linear: k must be positive and increasing");
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    std::vector<double> buf(3 * (size_t)n_k);
    const double twopi2 = 2.0 * M_PI * M_PI;
    for (int i = 0; i < n_k; ++i) {
        buf[i] = k[i];
        buf[n_k + i] = plin_z0[i];
        buf[2 * (size_t)n_k + i] = plin_z0[i] * (k[i] * k[i] * k[i]) / twopi2 * w_lnk[i];   // Delta0^2 * simps weight
    }
    // numpy-order Simpson data (oracle/simps121.py `_composite` with x given): per-panel constants of both panel sets
    std::vector<double> x(n_k), h(n_k - 1);
    for (int i = 0; i < n_k; ++i) x[i] = lnk ? lnk[i] : log(k[i]);
    for (int i = 0; i + 1 < n_k; ++i) { volatile double d = x[i + 1] - x[i]; h[i] = d; }
    const bool odd = n_k % 2 == 1;
    const int npan0 = odd ? (n_k - 1) / 2 : (n_k - 2) / 2, npan1 = odd ? 0 : (n_k - 2) / 2;
    std::vector<double> sp((size_t)n_k + 5 * (size_t)(npan0 + npan1));
    for (int i = 0; i < n_k; ++i) sp[i] = plin_z0[i] * (k[i] * k[i] * k[i]) / twopi2;       // Delta0^2
    for (int p = 0; p < npan0 + npan1; ++p) {
        const int i = p < npan0 ? 2 * p : 2 * (p - npan0) + 1;
        volatile double h0 = h[i], h1 = h[i + 1];
        volatile double hsum = h0 + h1, hprod = h0 * h1, q = h0 / h1;
        volatile double inv = 1.0 / q;
        volatile double c0 = 2 - inv, c2 = 2 - q, s6 = hsum / 6.0;
        double* o = &sp[(size_t)n_k + 5 * (size_t)p];
        o[0] = c0; o[1] = hsum; o[2] = hprod; o[3] = c2; o[4] = s6;
    }
    if (c->pk_lnk.reserve(buf.size() * 8) || c->pk_simps.reserve(sp.size() * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->pk_lnk.p, buf.data(), buf.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->pk_simps.p, sp.data(), sp.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaStreamSynchronize(c->stream));
    c->pk_npan0 = npan0; c->pk_npan1 = npan1;
    c->pk_first_dx = h[0]; c->pk_last_dx = h[n_k - 2];
    if (int rc = upload_rdk(c, n_k, k)) return rc;
    c->n_k = n_k;
    c->drop_pk_tables();
    return LBG_OK;
}

extern "C" int lbg_upload_pk_linear(lbg_ctx* c, int n_k, const double* k, const double* plin_z0, const double* w_lnk) {
    return lbg_upload_pk_linear2(c, n_k, k, plin_z0, w_lnk, nullptr);
}

extern "C" int lbg_halofit_sigma2(lbg_ctx* c, int n_R, const double* R, double* out_sigma2) {
    LBG_REQUIRE(c && R && out_sigma2 && n_R > 0, "lbg_halofit_sigma2: bad argument");
    if (c->n_k <= 0) { set_error("lbg_halofit_sigma2: call lbg_upload_pk_linear first"); return LBG_ESTATE; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    if (c->s0.reserve((size_t)n_R * 8) || c->s1.reserve((size_t)n_R * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, R, (size_t)n_R * 8, cudaMemcpyHostToDevice, c->stream));
    const double* kd = c->pk_lnk.as<double>();
    KernelTimer t(c);
    sigma2_kernel<<<n_R, 256, 0, c->stream>>>(c->n_k, kd, kd + 2 * (size_t)c->n_k, c->s0.as<double>(), c->s1.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_sigma2, c->s1.p, (size_t)n_R * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

extern "C" int lbg_pk_build_table(lbg_ctx* c, int n_rows, const double* halo_z /*[n_rows][8]*/, double fnu, double om0,
                                  double w, int takahashi, double* out_pnl /*[n_rows][n_k] or NULL*/) {
    LBG_REQUIRE(c && halo_z && n_rows > 0, "lbg_pk_build_table: bad argument");
    if (c->n_k <= 0) { set_error("lbg_pk_build_table: call lbg_upload_pk_linear first"); return LBG_ESTATE; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const int n_k = c->n_k;
    if (c->s5.reserve((size_t)n_rows * sizeof(HaloZ)) || c->pk_pnl.reserve((size_t)n_rows * n_k * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s5.p, halo_z, (size_t)n_rows * sizeof(HaloZ), cudaMemcpyHostToDevice, c->stream));
    HaloConst hc{fnu, om0, w, takahashi};
    const double* kd = c->pk_lnk.as<double>();
    KernelTimer t(c);
    pnl_kernel<<<dim3((n_k + 127) / 128, n_rows), 128, 0, c->stream>>>(n_rows, n_k, kd, kd + n_k, c->s5.as<HaloZ>(), hc, c->pk_pnl.as<double>());
    t.launched();
    if (out_pnl) LBG_CUDA(cudaMemcpyAsync(out_pnl, c->pk_pnl.p, (size_t)n_rows * n_k * 8, cudaMemcpyDeviceToHost, c->stream));
    int rc = t.finish();
    if (rc == LBG_OK) c->pnl_nz = n_rows;
    return rc;
}

extern "C" int lbg_pk_select_slot(lbg_ctx* c, int slot, int* out_rows) {
    LBG_REQUIRE(c, "lbg_pk_select_slot: ctx is NULL");
    LBG_REQUIRE(slot >= 0 && slot < lbg_ctx::PK_SLOTS, "lbg_pk_select_slot: slot %d outside [0, %d)", slot, lbg_ctx::PK_SLOTS);
    std::lock_guard<std::mutex> g(c->mu);
    if (slot != c->pk_cur) {                       // park the current table, bring the requested one in (pointer swaps, no GPU work)
        std::swap(c->pk_pnl, c->pk_store[c->pk_cur]);
        std::swap(c->pnl_nz, c->store_nz[c->pk_cur]);
        std::swap(c->pk_pnl, c->pk_store[slot]);
        std::swap(c->pnl_nz, c->store_nz[slot]);
        c->pk_cur = slot;
    }
    if (out_rows) *out_rows = c->pnl_nz;
    return LBG_OK;
}

// rows of a finished Limber batch moved into a kept spectra set: set row dest[b] <- batch row b (dest[b] < 0: dropped)
__global__ void keep_rows_kernel(int n_batch, int n_ell, const int* __restrict__ dest, const double* __restrict__ src, double* __restrict__ set) {
    const size_t n = (size_t)n_batch * n_ell, t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    const int b = (int)(t / n_ell), d = dest[b];
    if (d >= 0) set[(size_t)d * n_ell + (t - (size_t)b * n_ell)] = src[t];
}

// out_cl != NULL: the spectra go back to the host (lbg_limber).  set >= 0: they stay on the device in spectra set `set`, a
// [set_rows][n_ell] matrix, batch row b in set row dest_row[b] (lbg_limber_keep); rows of the set no batch entry names keep their contents.
static int limber_impl(lbg_ctx* c, int n_batch, int n_ell, const double* ell, int n_z, const double* chi, const double* weight,
                       const int* prow, int oob, double* out_cl, int set, int set_rows, const int* dest_row) {
    LBG_REQUIRE(c && ell && chi && weight && prow && (out_cl || set >= 0), "lbg_limber: NULL argument");
    LBG_REQUIRE(n_batch > 0 && n_ell >= 0 && n_z > 0, "lbg_limber: bad sizes");
    if (set >= 0) {
        LBG_REQUIRE(set < lbg_ctx::SPEC_SETS, "lbg_limber_keep: set %d outside [0, %d)", set, lbg_ctx::SPEC_SETS);
        LBG_REQUIRE(dest_row && set_rows > 0, "lbg_limber_keep: no destination rows");
        for (int b = 0; b < n_batch; ++b) LBG_REQUIRE(dest_row[b] < set_rows, "lbg_limber_keep: dest_row[%d]=%d outside the %d-row set", b, dest_row[b], set_rows);
    }
    std::lock_guard<std::mutex> g(c->mu);     // the table state is read under the lock: another thread may be swapping slots
    if (c->pnl_nz <= 0) { set_error("lbg_limber: call lbg_pk_build_table first"); return LBG_ESTATE; }
    for (int i = 0; i < n_z; ++i) LBG_REQUIRE(prow[i] >= 0 && prow[i] < c->pnl_nz, "lbg_limber: prow[%d]=%d outside the %d-row table", i, prow[i], c->pnl_nz);
    if (n_ell == 0) return LBG_OK;
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t bw = (size_t)n_batch * n_z * 8, bo = (size_t)n_batch * n_ell * 8;
    const int nbk = n_batch == 1 ? 1 : LIMBER_BT, n_bb = (n_batch + nbk - 1) / nbk;
    // every input of the call is packed into ONE pinned staging block and crosses PCIe in one copy (five pageable copies cost more
    // than the kernel): [ell | chi | RN(1/chi) | weight | prow (int) | live (bytes)], each piece 256-byte aligned
    auto up = [](size_t v) { return (v + 255) & ~(size_t)255; };
    const size_t o_ell = 0, o_zc = up((size_t)n_ell * 8), o_w = o_zc + up((size_t)n_z * 16), o_prow = o_w + up(bw),
                 o_live = o_prow + up((size_t)n_z * 4), total = o_live + up((size_t)n_bb * n_z);
    if (c->s0.reserve(total) || c->s4.reserve(bo) || c->pin_in.reserve(total) || c->pin_out.reserve(bo)) return LBG_ENOMEM;
    const int n_split = limber_splits(n_z), z_per_split = (n_z + n_split - 1) / n_split;
    if (n_split > 1 && c->s5.reserve(bo * n_split)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    unsigned char* hin = c->pin_in.as<unsigned char>();
    memcpy(hin + o_ell, ell, (size_t)n_ell * 8);
    double* zc = reinterpret_cast<double*>(hin + o_zc);       // chi | RN(1 / chi) for div_rn
    for (int i = 0; i < n_z; ++i) { zc[i] = chi[i]; zc[(size_t)n_z + i] = 1.0 / chi[i]; }
    memcpy(hin + o_w, weight, bw);
    memcpy(hin + o_prow, prow, (size_t)n_z * 4);
    // a z node none of a block's batch entries weights is skipped by that block.  Deliberate deviation from the reference's
    // simps(prefactor * result): a NaN / Inf P(k,z) row under an all-zero weight gives 0 here, NaN there
    unsigned char* lv = hin + o_live;
    for (int bb = 0; bb < n_bb; ++bb)
        for (int i = 0; i < n_z; ++i) {
            bool any = false;
            for (int q = bb * nbk; q < std::min(n_batch, (bb + 1) * nbk); ++q) any = any || (weight[(size_t)q * n_z + i] != 0.0);
            lv[(size_t)bb * n_z + i] = any;
        }
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, hin, total, cudaMemcpyHostToDevice, st));
    const unsigned char* din = c->s0.as<unsigned char>();
    const double* d_ell = reinterpret_cast<const double*>(din + o_ell);
    const double* d_zc = reinterpret_cast<const double*>(din + o_zc);
    const double* d_w = reinterpret_cast<const double*>(din + o_w);
    const int* d_prow = reinterpret_cast<const int*>(din + o_prow);
    const unsigned char* live = din + o_live;
    KernelTimer t(c);
    double* part = n_split > 1 ? c->s5.as<double>() : c->s4.as<double>();
    const dim3 grid((n_ell + 127) / 128, n_bb, n_split);
#define LBG_LIMBER(NB) limber_kernel<NB><<<grid, 128, 0, st>>>(n_batch, n_ell, n_z, c->n_k, z_per_split, d_ell, d_zc, d_w, live, d_prow, \
        c->pk_lnk.as<double>(), c->pk_rdk.as<double>(), c->pk_pnl.as<double>(), oob, part)
    if (nbk == 1) LBG_LIMBER(1); else LBG_LIMBER(LIMBER_BT);
#undef LBG_LIMBER
    t.launched();
    if (n_split > 1) {
        const size_t n = (size_t)n_batch * n_ell;
        limber_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, n_split, part, c->s4.as<double>());
        t.launched();
    }
    if (set >= 0) {
        // a set keeps its shape until a call names another one: the first call of a new shape starts from zeros
        if (c->set_rows[set] != set_rows || c->set_n_ell[set] != n_ell) {
            if (c->spec_set[set].reserve((size_t)set_rows * n_ell * 8)) return LBG_ENOMEM;
            LBG_CUDA(cudaMemsetAsync(c->spec_set[set].p, 0, (size_t)set_rows * n_ell * 8, st));
            c->set_rows[set] = set_rows; c->set_n_ell[set] = n_ell;
        }
        if (c->s3.reserve((size_t)n_batch * 4)) return LBG_ENOMEM;
        int* hd = reinterpret_cast<int*>(c->pin_out.p);          // pin_out holds >= n_batch * n_ell * 8 bytes and is otherwise unused here
        memcpy(hd, dest_row, (size_t)n_batch * 4);
        LBG_CUDA(cudaMemcpyAsync(c->s3.p, hd, (size_t)n_batch * 4, cudaMemcpyHostToDevice, st));
        const size_t n = (size_t)n_batch * n_ell;
        keep_rows_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n_batch, n_ell, c->s3.as<int>(), c->s4.as<double>(), c->spec_set[set].as<double>());
        t.launched();
    }
    if (out_cl) LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, c->s4.p, bo, cudaMemcpyDeviceToHost, st));
    const int rc = t.finish();
    if (rc == LBG_OK && out_cl) memcpy(out_cl, c->pin_out.p, bo);
    return rc;
}

extern "C" int lbg_limber(lbg_ctx* c, int n_batch, int n_ell, const double* ell, int n_z, const double* chi,
                          const double* weight /*[n_batch][n_z]*/, const int* prow /*[n_z]*/, int oob,
                          double* out_cl /*[n_batch][n_ell]*/) {
    LBG_REQUIRE(out_cl, "lbg_limber: NULL argument");
    return limber_impl(c, n_batch, n_ell, ell, n_z, chi, weight, prow, oob, out_cl, -1, 0, nullptr);
}

extern "C" int lbg_limber_keep(lbg_ctx* c, int n_batch, int n_ell, const double* ell, int n_z, const double* chi,
                               const double* weight, const int* prow, int oob, int set, int set_rows, const int* dest_row) {
    LBG_REQUIRE(set >= 0, "lbg_limber_keep: set %d outside [0, %d)", set, lbg_ctx::SPEC_SETS);
    return limber_impl(c, n_batch, n_ell, ell, n_z, chi, weight, prow, oob, nullptr, set, set_rows, dest_row);
}

// Rows of a kept set filled without a Limber sum (noise): row r = row_const[r] at every multipole, then rows curve_row[i] replaced
// by the caller's curves[i][n_ell].
__global__ void set_fill_kernel(int n_rows, int n_ell, const double* __restrict__ row_const, double* __restrict__ set) {
    const size_t n = (size_t)n_rows * n_ell, t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) set[t] = row_const[t / n_ell];
}

extern "C" int lbg_spectra_set_fill(lbg_ctx* c, int set, int set_rows, int n_ell, const double* row_const, int n_curves,
                                    const int* curve_row, const double* curves) {
    LBG_REQUIRE(c && row_const && (n_curves == 0 || (curve_row && curves)), "lbg_spectra_set_fill: NULL argument");
    LBG_REQUIRE(set >= 0 && set < lbg_ctx::SPEC_SETS, "lbg_spectra_set_fill: set %d outside [0, %d)", set, lbg_ctx::SPEC_SETS);
    LBG_REQUIRE(set_rows > 0 && n_ell >= 0 && n_curves >= 0, "lbg_spectra_set_fill: bad sizes");
    for (int i = 0; i < n_curves; ++i) LBG_REQUIRE(curve_row[i] >= 0 && curve_row[i] < set_rows, "lbg_spectra_set_fill: curve_row[%d]=%d outside the %d-row set", i, curve_row[i], set_rows);
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    c->set_rows[set] = set_rows; c->set_n_ell[set] = n_ell;
    if (n_ell == 0) return LBG_OK;
    const size_t bc = ((size_t)set_rows * 8 + 255) & ~(size_t)255, total = bc + (size_t)n_curves * n_ell * 8;
    if (c->spec_set[set].reserve((size_t)set_rows * n_ell * 8) || c->s0.reserve(total) || c->pin_in.reserve(total)) return LBG_ENOMEM;
    unsigned char* hin = c->pin_in.as<unsigned char>();
    memcpy(hin, row_const, (size_t)set_rows * 8);
    if (n_curves) memcpy(hin + bc, curves, (size_t)n_curves * n_ell * 8);
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, hin, total, cudaMemcpyHostToDevice, st));
    KernelTimer t(c);
    const size_t n = (size_t)set_rows * n_ell;
    set_fill_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(set_rows, n_ell, c->s0.as<double>(), c->spec_set[set].as<double>());
    t.launched();
    for (int i = 0; i < n_curves; ++i)
        LBG_CUDA(cudaMemcpyAsync(c->spec_set[set].as<double>() + (size_t)curve_row[i] * n_ell, c->s0.as<unsigned char>() + bc + (size_t)i * n_ell * 8,
                                 (size_t)n_ell * 8, cudaMemcpyDeviceToDevice, st));
    return t.finish();
}

extern "C" int lbg_spectra_set_get(lbg_ctx* c, int set, int row0, int n_rows, int n_ell, double* out) {
    LBG_REQUIRE(c && out, "lbg_spectra_set_get: NULL argument");
    LBG_REQUIRE(set >= 0 && set < lbg_ctx::SPEC_SETS, "lbg_spectra_set_get: set %d outside [0, %d)", set, lbg_ctx::SPEC_SETS);
    std::lock_guard<std::mutex> g(c->mu);
    if (c->set_rows[set] <= 0 || c->set_n_ell[set] != n_ell) { set_error("lbg_spectra_set_get: set %d holds %d rows of %d multipoles, asked for %d multipoles", set, c->set_rows[set], c->set_n_ell[set], n_ell); return LBG_ESTATE; }
    LBG_REQUIRE(row0 >= 0 && n_rows >= 0 && row0 + n_rows <= c->set_rows[set], "lbg_spectra_set_get: rows [%d, %d) outside the %d-row set", row0, row0 + n_rows, c->set_rows[set]);
    if (n_rows == 0 || n_ell == 0) return LBG_OK;
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t b = (size_t)n_rows * n_ell * 8;
    if (c->pin_out.reserve(b)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, c->spec_set[set].as<double>() + (size_t)row0 * n_ell, b, cudaMemcpyDeviceToHost, c->stream));
    LBG_CUDA(cudaStreamSynchronize(c->stream));
    memcpy(out, c->pin_out.p, b);
    return LBG_OK;
}

extern "C" int lbg_pk_eval(lbg_ctx* c, int row, int n, const double* x, int oob, double* out) {
    LBG_REQUIRE(c && x && out, "lbg_pk_eval: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    if (c->pnl_nz <= 0) { set_error("lbg_pk_eval: call lbg_pk_build_table first"); return LBG_ESTATE; }
    LBG_REQUIRE(row >= 0 && row < c->pnl_nz, "lbg_pk_eval: row %d outside the %d-row table", row, c->pnl_nz);
    if (n <= 0) return LBG_OK;
    LBG_CUDA(cudaSetDevice(c->device));
    if (c->s0.reserve((size_t)n * 8) || c->s1.reserve((size_t)n * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, x, (size_t)n * 8, cudaMemcpyHostToDevice, c->stream));
    KernelTimer t(c);
    pk_eval_kernel<<<(n + 127) / 128, 128, 0, c->stream>>>(n, c->n_k, c->pk_lnk.as<double>(), c->pk_pnl.as<double>() + (size_t)row * c->n_k,
                                                          c->s0.as<double>(), oob, c->s1.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out, c->s1.p, (size_t)n * 8, cudaMemcpyDeviceToHost, c->stream));
    return t.finish();
}

// Caller-supplied P(k) rows on their own k grid (the `kp = (k, P)` tables of modi/lensing.py:127-167): replaces the
// device table; a later lbg_pk_build_table needs lbg_upload_pk_linear again.
extern "C" int lbg_pk_upload_rows(lbg_ctx* c, int n_k, const double* k, int n_rows, const double* rows) {
    LBG_REQUIRE(c && k && rows, "lbg_pk_upload_rows: NULL argument");
    LBG_REQUIRE(n_k >= 2 && n_k <= 6000 && n_rows > 0, "lbg_pk_upload_rows: bad sizes n_k=%d n_rows=%d (n_k <= 6000: the k grid is staged in shared memory)", n_k, n_rows);
    for (int i = 1; i < n_k; ++i) LBG_REQUIRE(k[i] > k[i - 1], "lbg_pk_upload_rows: k must be increasing");
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    std::vector<double> buf(3 * (size_t)n_k, 0.0);
    memcpy(buf.data(), k, (size_t)n_k * 8);
    if (c->pk_lnk.reserve(buf.size() * 8) || c->pk_pnl.reserve((size_t)n_rows * n_k * 8)) return LBG_ENOMEM;
    LBG_CUDA(cudaMemcpyAsync(c->pk_lnk.p, buf.data(), buf.size() * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaMemcpyAsync(c->pk_pnl.p, rows, (size_t)n_rows * n_k * 8, cudaMemcpyHostToDevice, c->stream));
    LBG_CUDA(cudaStreamSynchronize(c->stream));
    if (int rc = upload_rdk(c, n_k, k)) return rc;
    c->n_k = n_k;
    c->pk_npan0 = c->pk_npan1 = 0;
    c->drop_pk_tables();
    c->pnl_nz = n_rows;
    return LBG_OK;
}


// One Limber sum on device-resident inputs (the body of lbg_limber without its copies); weights / live / partial are the caller's.
static int limber_launch_dev(lbg_ctx* c, cudaStream_t st, int n_batch, int n_ell, const double* d_ell, int n_z, const double* d_zc,
                             const double* d_w, unsigned char* d_live, const int* d_prow, const double* d_table, int oob,
                             double* d_partial, double* d_out, int* launches) {
    const int nbk = n_batch == 1 ? 1 : LIMBER_BT, n_bb = (n_batch + nbk - 1) / nbk;
    const int n_split = limber_splits(n_z), z_per_split = (n_z + n_split - 1) / n_split;
    limber_live_kernel<<<dim3((n_z + 127) / 128, n_bb), 128, 0, st>>>(n_batch, nbk, n_z, d_w, d_live);
    double* part = n_split > 1 ? d_partial : d_out;
    const dim3 grid((n_ell + 127) / 128, n_bb, n_split);
#define LBG_LIMBER(NB) limber_kernel<NB><<<grid, 128, 0, st>>>(n_batch, n_ell, n_z, c->n_k, z_per_split, d_ell, d_zc, d_w, d_live, d_prow, \
        c->pk_lnk.as<double>(), c->pk_rdk.as<double>(), d_table, oob, part)
    if (nbk == 1) LBG_LIMBER(1); else LBG_LIMBER(LIMBER_BT);
#undef LBG_LIMBER
    *launches += 2;
    if (n_split > 1) {
        const size_t n = (size_t)n_batch * n_ell;
        limber_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, n_split, part, d_out);
        ++*launches;
    }
    return LBG_OK;
}

extern "C" int lbg_forecast_spectra(lbg_ctx* c, int n_ell, const double* ell, double zmin, double zmax, double h100,
                                    int n_gg, const double* z_gg, const double* chi_gg, const double* hc_gg, const double* w_gg, double dz_gg, int slot_gg,
                                    int n_kg, const double* z_kg, const double* chi_kg, const double* wk_kg, const double* w_kg, double dz_kg, int slot_kg,
                                    int n_kk, const double* chi_kk, const double* weight_kk, int slot_kk,
                                    int n_s, const double* pz_gg, const double* bz_gg, const double* pz_kg, const double* bz_kg, const double* norm,
                                    double* out_cgg, double* out_ckg, double* out_ckk) {
    LBG_REQUIRE(c && ell && z_gg && chi_gg && hc_gg && w_gg && z_kg && chi_kg && wk_kg && w_kg && pz_gg && bz_gg && pz_kg && bz_kg && norm &&
                out_cgg && out_ckg, "lbg_forecast_spectra: NULL argument");
    LBG_REQUIRE(n_ell >= 0 && n_gg > 0 && n_kg > 0 && n_s > 0 && n_kk >= 0, "lbg_forecast_spectra: bad sizes");
    LBG_REQUIRE(n_kk == 0 || (chi_kk && weight_kk && out_ckk), "lbg_forecast_spectra: C_kk grid given without its arrays / output");
    LBG_REQUIRE(h100 > 0 && dz_gg > 0 && dz_kg > 0, "lbg_forecast_spectra: h100 and the grid spacings must be positive");
    if (n_ell == 0) return LBG_OK;
    std::lock_guard<std::mutex> g(c->mu);
    if (c->n_k <= 0) { set_error("lbg_forecast_spectra: call lbg_upload_pk_linear first"); return LBG_ESTATE; }
    const int slots[3] = {slot_gg, slot_kg, slot_kk}, want[3] = {n_gg, n_kg, n_kk};
    const double* table[3] = {nullptr, nullptr, nullptr};
    for (int q = 0; q < 3; ++q) {
        if (want[q] == 0) continue;
        LBG_REQUIRE(slots[q] >= 0 && slots[q] < lbg_ctx::PK_SLOTS, "lbg_forecast_spectra: table slot %d outside [0, %d)", slots[q], lbg_ctx::PK_SLOTS);
        const bool cur = slots[q] == c->pk_cur;
        const int rows = cur ? c->pnl_nz : c->store_nz[slots[q]];
        if (rows != want[q]) { set_error("lbg_forecast_spectra: table slot %d holds %d rows, the grid has %d nodes (build it with lbg_pk_build_table)", slots[q], rows, want[q]); return LBG_ESTATE; }
        table[q] = cur ? c->pk_pnl.as<double>() : c->pk_store[slots[q]].as<double>();
    }
    LBG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    // ---- one pinned staging block in ------------------------------------------------------------------------------------------
    const size_t n_in = (size_t)n_ell + 2 * ((size_t)n_gg + n_kg + n_kk) + 3 * (size_t)n_gg + 3 * (size_t)n_kg + n_kk +
                        2 * (size_t)n_s * ((size_t)n_gg + n_kg) + n_s;
    const int n_zmax = std::max(n_gg, std::max(n_kg, n_kk));
    const int n_bb = (n_s + LIMBER_BT - 1) / LIMBER_BT;
    const size_t n_out = (size_t)(2 * n_s + (n_kk ? 1 : 0)) * n_ell;
    const size_t n_w = (size_t)n_s * ((size_t)n_gg + n_kg);
    const size_t part_sz = (size_t)LIMBER_SMAX * std::max(n_s, 1) * n_ell;
    if (c->s0.reserve(n_in * 8 + 64) || c->pin_in.reserve(n_in * 8 + 64) || c->s4.reserve(n_out * 8 + 64) || c->pin_out.reserve(n_out * 8 + 64) ||
        c->s1.reserve(n_w * 8 + 64) || c->s2.reserve((size_t)n_bb * n_zmax + 64) || c->s3.reserve((size_t)n_zmax * 4 + 64) || c->s5.reserve(part_sz * 8))
        return LBG_ENOMEM;
    double* h = c->pin_in.as<double>();
    size_t o = 0;
    auto put = [&](const double* src, size_t n) { size_t at = o; if (n) memcpy(h + o, src, n * 8); o += n; return at; };
    auto put_zc = [&](const double* chi, int n) { size_t at = o; for (int i = 0; i < n; ++i) { h[o + i] = chi[i]; h[o + n + i] = 1.0 / chi[i]; } o += 2 * (size_t)n; return at; };
    const size_t o_ell = put(ell, n_ell);
    const size_t o_zc_gg = put_zc(chi_gg, n_gg), o_zc_kg = put_zc(chi_kg, n_kg), o_zc_kk = put_zc(chi_kk, n_kk);
    const size_t o_hc = put(hc_gg, n_gg), o_wgg = put(w_gg, n_gg), o_zgg = put(z_gg, n_gg);
    const size_t o_wk = put(wk_kg, n_kg), o_wkg = put(w_kg, n_kg), o_zkg = put(z_kg, n_kg);
    const size_t o_wkk = put(weight_kk, n_kk);
    const size_t o_pgg = put(pz_gg, (size_t)n_s * n_gg), o_bgg = put(bz_gg, (size_t)n_s * n_gg);
    const size_t o_pkg = put(pz_kg, (size_t)n_s * n_kg), o_bkg = put(bz_kg, (size_t)n_s * n_kg);
    const size_t o_norm = put(norm, n_s);
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, h, o * 8, cudaMemcpyHostToDevice, st));
    const double* d = c->s0.as<double>();
    double* d_wgg = c->s1.as<double>();
    double* d_wkg = d_wgg + (size_t)n_s * n_gg;
    double* d_out = c->s4.as<double>();
    KernelTimer t(c);
    int launches = 0;
    iota_kernel<<<(n_zmax + 255) / 256, 256, 0, st>>>(n_zmax, c->s3.as<int>());
    WeightGrid G0{n_gg, d + o_zgg, d + o_zc_gg, d + o_hc, d + o_wgg, d + o_pgg, d + o_bgg, d_wgg, dz_gg};
    WeightGrid G1{n_kg, d + o_zkg, d + o_zc_kg, d + o_wk, d + o_wkg, d + o_pkg, d + o_bkg, d_wkg, dz_kg};
    forecast_weights_kernel<<<dim3(n_s, 2), 256, 0, st>>>(G0, G1, d + o_norm, zmin, zmax, h100);
    launches += 2;
    limber_launch_dev(c, st, n_s, n_ell, d + o_ell, n_gg, d + o_zc_gg, d_wgg, c->s2.as<unsigned char>(), c->s3.as<int>(), table[0], 0,
                      c->s5.as<double>(), d_out, &launches);
    limber_launch_dev(c, st, n_s, n_ell, d + o_ell, n_kg, d + o_zc_kg, d_wkg, c->s2.as<unsigned char>(), c->s3.as<int>(), table[1], 0,
                      c->s5.as<double>(), d_out + (size_t)n_s * n_ell, &launches);
    if (n_kk)
        limber_launch_dev(c, st, 1, n_ell, d + o_ell, n_kk, d + o_zc_kk, d + o_wkk, c->s2.as<unsigned char>(), c->s3.as<int>(), table[2], 0,
                          c->s5.as<double>(), d_out + (size_t)2 * n_s * n_ell, &launches);
    t.launched(launches);
    LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, d_out, n_out * 8, cudaMemcpyDeviceToHost, st));
    const int rc = t.finish();
    if (rc) return rc;
    const double* r = c->pin_out.as<double>();
    memcpy(out_cgg, r, (size_t)n_s * n_ell * 8);
    memcpy(out_ckg, r + (size_t)n_s * n_ell, (size_t)n_s * n_ell * 8);
    if (n_kk) memcpy(out_ckk, r + (size_t)2 * n_s * n_ell, (size_t)n_ell * 8);
    return LBG_OK;
}
