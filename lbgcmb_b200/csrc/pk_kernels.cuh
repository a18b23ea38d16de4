// pk_kernels.cuh -- the two matter-power kernels shared by limber.cu (explicit-row table builds) and halofit.cu (device-side spectral
// parameters): sigma^2(R) integrals (modi/halofit.py:36-44) and the Takahashi / Smith formula on every table node (:139-212).
#pragma once
#include "common.cuh"
#include <cmath>

namespace lbg {

// sigma0^2(R) = sum_i w_i Delta0^2(k_i) exp(-(k_i R)^2): one block per R.  w = scipy-1.2.1 simps weights in ln k.
static __global__ void __launch_bounds__(256)
sigma2_kernel(int n_k, const double* __restrict__ k, const double* __restrict__ d2w /*Delta0^2 * w*/,
              const double* __restrict__ R, double* __restrict__ out) {
    __shared__ double sm[256];
    const double r = R[blockIdx.x];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n_k; i += 256) {
        const double kr = k[i] * r;
        acc += d2w[i] * exp(-(kr * kr));
    }
    sm[threadIdx.x] = acc;
    __syncthreads();
    for (int h = 128; h > 0; h >>= 1) {
        if (threadIdx.x < h) sm[threadIdx.x] += sm[threadIdx.x + h];
        __syncthreads();
    }
    if (threadIdx.x == 0) out[blockIdx.x] = sm[0];
}

// The same integral summed EXACTLY as the reference sums it (modi/halofit.py:41-42: scipy-1.2.1 `simps(y, ln k)` = numpy products per
// three-point panel + np.sum, whose pairwise order is reproduced below).  Needed where ln sigma^2(R) feeds HALOFIT's curvature: the
// second derivative of the interpolating spline multiplies the ~1e-16 difference between two summation orders by ~1e4 / C.
// Block per R: threads fill y_i = Delta0^2 exp(-(k_i R)^2) and the panel terms in shared memory and add them in numpy's order (below).
struct SimpsPanel { double c0, hsum, hprod, c2, s6; };   // 2 - 1/(h0/h1), h0 + h1, h0 h1, 2 - h0/h1, (h0 + h1)/6 of one panel

__device__ inline double np_pairwise_sum(const double* a, int n) {       // numpy's pairwise_sum for contiguous float64
    if (n < 8) {
        double res = 0.0;
        for (int i = 0; i < n; ++i) res = __dadd_rn(res, a[i]);
        return res;
    }
    if (n <= 128) {
        double r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int i = 8;
        for (; i < n - (n % 8); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
        }
        double res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])), __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
        for (; i < n; ++i) res = __dadd_rn(res, a[i]);
        return res;
    }
    int n2 = n / 2;
    n2 -= n2 % 8;
    return __dadd_rn(np_pairwise_sum(a, n2), np_pairwise_sum(a + n2, n - n2));
}

// numpy's pairwise sum is a binary recursion over halves down to leaves of <= 128 elements.  The tree (pre-order: a parent before its
// children) is laid out by one thread; the leaves are then independent sums that any thread can take, and the inner nodes are
// combined child + child in reverse index order -- the additions numpy makes, in its order.
struct PwNode { int off, n, left, right; };      // left < 0: leaf
constexpr int PW_MAX_NODES = 96;                 // 2 trees of up to ~1000 panels: 2 x (2 x 16 - 1) nodes; bound checked on the host

__device__ inline int pw_build(PwNode* nodes, int* cnt, int off, int n) {
    const int me = (*cnt)++;
    nodes[me].off = off; nodes[me].n = n; nodes[me].left = -1; nodes[me].right = -1;
    if (n > 128) {
        int n2 = n / 2;
        n2 -= n2 % 8;
        const int l = pw_build(nodes, cnt, off, n2);
        const int r = pw_build(nodes, cnt, off + n2, n - n2);
        nodes[me].left = l; nodes[me].right = r;
    }
    return me;
}

static __global__ void __launch_bounds__(256, 4)
sigma2_np_kernel(int n, const double* __restrict__ k, const double* __restrict__ delta0, const SimpsPanel* __restrict__ pan,
                 int npan0, int npan1, double first_dx, double last_dx, const double* __restrict__ R, double* __restrict__ out) {
    extern __shared__ double sh[];
    double* y = sh;                       // [n]
    double* tmp = sh + n;                 // [npan0 + npan1]
    __shared__ PwNode nodes[PW_MAX_NODES];
    __shared__ double val[PW_MAX_NODES];
    __shared__ int roots[3];              // root of the first tree, root of the second, node count
    const double r = R[blockIdx.x];
    if (threadIdx.x == 255) {             // the summation trees, laid out while the other threads fill y
        int cnt = 0;
        roots[0] = pw_build(nodes, &cnt, 0, npan0);
        roots[1] = npan1 > 0 ? pw_build(nodes, &cnt, npan0, npan1) : -1;
        roots[2] = cnt;
    }
    for (int i = threadIdx.x; i < n; i += 256) {
        const double kr = __dmul_rn(k[i], r);
        y[i] = __dmul_rn(delta0[i], exp(-__dmul_rn(kr, kr)));
    }
    __syncthreads();
    for (int p = threadIdx.x; p < npan0 + npan1; p += 256) {
        const int i = p < npan0 ? 2 * p : 2 * (p - npan0) + 1;            // first sample of the panel
        const SimpsPanel P = pan[p];
        const double a = __dmul_rn(y[i], P.c0);
        const double b = __ddiv_rn(__dmul_rn(__dmul_rn(y[i + 1], P.hsum), P.hsum), P.hprod);
        const double c = __dmul_rn(y[i + 2], P.c2);
        tmp[p] = __dmul_rn(P.s6, __dadd_rn(__dadd_rn(a, b), c));
    }
    __syncthreads();
    const int cnt = roots[2];
    // one leaf per warp-strided thread: leaves are <= 128 elements, 8 independent accumulators each
    for (int q = threadIdx.x; q < cnt; q += 256)
        if (nodes[q].left < 0) val[q] = np_pairwise_sum(tmp + nodes[q].off, nodes[q].n);
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = cnt - 1; q >= 0; --q)
            if (nodes[q].left >= 0) val[q] = __dadd_rn(val[nodes[q].left], val[nodes[q].right]);
        double res = val[roots[0]];
        if (npan1 > 0) {                  // even number of samples: scipy-1.2.1 even='avg'
            double v2 = __dmul_rn(__dmul_rn(0.5, last_dx), __dadd_rn(y[n - 1], y[n - 2]));
            v2 = __dadd_rn(v2, __dmul_rn(__dmul_rn(0.5, first_dx), __dadd_rn(y[1], y[0])));
            res = __dadd_rn(res, val[roots[1]]);
            res = __dadd_rn(__ddiv_rn(res, 2.0), __ddiv_rn(v2, 2.0));
        }
        out[blockIdx.x] = res;
    }
}

struct HaloZ {           // per-redshift HALOFIT inputs (modi/halofit.py:139-199)
    double D2;           // growth^2 applied to the z=0 table (modi/cosmology.py:160,175)
    double rknl, neff, ncur, omz, ovz;
    double use_halofit;  // 0 -> linear row (z > 4, modi/cosmology.py:171-172)
    double pad;
};

struct HaloConst { double fnu, om0, w; int takahashi; };

// One thread per (z, k) table node: P_nl(k, z) [(Mpc/h)^3] from Delta_lin^2 = P0 k^3 D^2 / (2 pi^2).  A block works on ONE redshift
// (blockIdx.y), and ten of the fourteen pow() calls of the formula depend on the redshift only: the first warp takes them in one
// call -- lane i its own (base, exponent) pair, no divergence -- and leaves them in shared memory; a node then costs four pow + one exp.
static __global__ void pnl_kernel(int n_z, int n_k, const double* __restrict__ k, const double* __restrict__ p0,
                           const HaloZ* __restrict__ hz, HaloConst hc, double* __restrict__ out) {
    const int ik = blockIdx.x * blockDim.x + threadIdx.x;
    const int iz = blockIdx.y;
    __shared__ double pw[16];
    const HaloZ h = hz[iz];
    const double neff = h.neff, rncur = h.ncur, w = hc.w, fnu = hc.fnu;
    const double n2 = neff * neff, n3 = n2 * neff, n4 = n2 * n2;
    const bool nl = h.use_halofit != 0.0;                                 // block-uniform
    if (nl && threadIdx.x < 11) {
        double base = 10.0, ex = 0.0;
        if (hc.takahashi) {                                               // modi/halofit.py:152-166
            switch (threadIdx.x) {
                case 0: ex = 1.5222 + 2.8553 * neff + 2.3706 * n2 + 0.9903 * n3 + 0.2250 * n4 - 0.6038 * rncur + 0.1749 * h.ovz * (1 + w); break;
                case 1: ex = -0.5642 + 0.5864 * neff + 0.5716 * n2 - 1.5474 * rncur + 0.2279 * h.ovz * (1 + w); break;
                case 2: ex = 0.3698 + 2.0404 * neff + 0.8161 * n2 + 0.5869 * rncur; break;
                case 3: ex = 5.2105 + 3.6902 * neff; break;
                default: break;
            }
        } else {                                                          // modi/halofit.py:168-180 (Smith+03)
            switch (threadIdx.x) {
                case 0: ex = 1.4861 + 1.8369 * neff + 1.6762 * n2 + 0.7940 * n3 + 0.1670 * n4 - 0.6206 * rncur; break;
                case 1: ex = 0.9463 + 0.9466 * neff + 0.3084 * n2 - 0.94 * rncur; break;
                case 2: ex = -0.2807 + 0.6669 * neff + 0.3214 * n2 - 0.0793 * rncur; break;
                case 3: ex = 0.9589 + 1.2857 * neff; break;
                case 4: ex = -3.5442 + 0.1908 * neff; break;
                default: break;
            }
        }
        if (threadIdx.x >= 5) {                                           // modi/halofit.py:183-199: f1a f2a f3a f1b f2b f3b
            const double fe[6] = {-0.0732, -0.1423, 0.0725, -0.0307, -0.0585, 0.0743};
            base = h.omz;
            ex = fe[threadIdx.x - 5];
        }
        pw[threadIdx.x] = pow(base, ex);                                  // the exponents were plain arithmetic: one pow for the warp
    }
    __syncthreads();
    if (ik >= n_k) return;
    const double kk = k[ik];
    const double twopi2 = 2.0 * 3.14159265358979323846 * 3.14159265358979323846;
    const double plin_dim = p0[ik] * (kk * kk * kk) / twopi2 * h.D2;     // modi/cosmology.py:174-175
    double res = plin_dim;
    if (nl && kk > 0.005) {                                               // modi/halofit.py:139
        const double a = pw[0], b = pw[1], c = pw[2], xnu = pw[3];
        double gam, alpha, beta, xmu;
        if (hc.takahashi) {
            gam = 0.1971 - 0.0843 * neff + 0.8460 * rncur;
            alpha = fabs(6.0835 + 1.3373 * neff - 0.1959 * n2 - 5.5274 * rncur);
            beta = 2.0379 - 0.7354 * neff + 0.3157 * n2 + 1.2490 * n3 + 0.3980 * n4 - 0.1682 * rncur + fnu * (1.081 + 0.395 * n2);
            xmu = 0.0;
        } else {
            gam = 0.8649 + 0.2989 * neff + 0.1631 * rncur;
            alpha = fabs(1.3884 + 0.3700 * neff - 0.1452 * n2);
            beta = 0.8291 + 0.9854 * neff + 0.3401 * n2;
            xmu = pw[4];
        }
        double f1 = 1.0, f2 = 1.0, f3 = 1.0;
        if (fabs(1 - h.omz) > 0.01) {                                     // modi/halofit.py:183-199
            const double f1a = pw[5], f2a = pw[6], f3a = pw[7], f1b = pw[8], f2b = pw[9], f3b = pw[10];
            const double frac = h.ovz / (1 - h.omz);
            if (hc.takahashi) { f1 = f1b; f2 = f2b; f3 = f3b; }
            else { f1 = frac * f1b + (1 - frac) * f1a; f2 = frac * f2b + (1 - frac) * f2a; f3 = frac * f3b + (1 - frac) * f3a; }
        }
        const double y = kk / h.rknl;                                     // modi/halofit.py:201-212
        double ph = a * pow(y, f1 * 3) / (1 + b * pow(y, f2) + pow(f3 * c * y, 3 - gam));
        ph = ph / (1 + xmu / y + xnu / (y * y)) * (1 + fnu * (0.977 - 18.015 * (hc.om0 - 0.3)));
        const double plinaa = plin_dim * (1 + fnu * 47.48 * kk * kk / (1 + 1.5 * kk * kk));
        const double pq = plin_dim * pow(1 + plinaa, beta) / (1 + plinaa * alpha) * exp(-y / 4.0 - y * y / 8.0);
        res = pq + ph;
    }
    out[(size_t)iz * n_k + ik] = res * twopi2 / (kk * kk * kk);           // modi/cosmology.py:179
}

}  // namespace lbg
