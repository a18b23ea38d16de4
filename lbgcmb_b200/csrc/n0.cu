// n0.cu -- Hu-Okamoto flat-sky N0 reconstruction noise, root flavour (replaces cmb/Nkk.py:54-172).
//
// Work decomposition (B200: 148 SMs, FP64-pipe bound, tables L2-resident):
//   thread  <-> one (L, phi_j) pair; it marches over a chunk of the l1 = |u| grid.
//   Every quantity that depends on l1 only (unlensed/total spectra at u, their reciprocals, weight*l1)
//   is precomputed once per noise configuration into a 128-byte "u row" that the whole block reads
//   at the same address (broadcast).  The v-side spectra are eight cubic splines evaluated from a
//   knot-major table of uniform B-spline coefficients (64 B per knot, four consecutive knots per node).
//   Partial sums go to a scratch buffer and are reduced in a fixed order (deterministic, no atomics).
#include "common.cuh"
#include <cmath>

namespace lbg {

struct __align__(128) URow {
    double l, wl;                 // node, quadrature weight * l   (cmb/Nkk.py:164 `result *= l`)
    double uT, uX, uE, uB;        // unlensed TT, TE, EE, BB at l
    double T1, E1, X1;            // lensed + noise TT, EE ; lensed TE  (prep_camb.py:128-137)
    double rT1h, rE1h, rE1, rB1h; // 0.5/T1, 0.5/E1, 1/E1, 0.5/B1
    double flags;                 // bit0 T-terms active, bit2 P-terms active, bit4 row handled by n0_special_kernel
    double pad0, pad1;
};

struct N0Params {
    int n_L, n_phi, n_l1, n_chunk, rows_per_chunk, n_ell;
    int soa_stride;               // > 0: `rows` of n0_node is the spectrum-major table double[8][soa_stride]
    unsigned terms;
    double nT, nP, a;             // N_TT(l) = nT exp(a l (l+1)), N_EE = N_BB = nP exp(.)  (cmb/DetectorNoise.py:22-26)
    double lmax_T, lmax_P, lmin_v;
    double ebfac;
    int fold;                     // phi nodes are mirror pairs: phi[] holds [6][n_phi] cos,sin,w,cos_m,sin_m,frac_j
    unsigned near_lo[3], near_span[3];   // hi-word windows of |v|^2 around lmin_v^2, lmax_T^2, lmax_P^2
};

__device__ __forceinline__ void bspline_weights(double t, double& w0, double& w1, double& w2, double& w3) {
    double um = 1.0 - t;
    double t2 = t * t, u2 = um * um;
    w0 = u2 * um * (1.0 / 6.0);
    w3 = t2 * t * (1.0 / 6.0);
    w1 = fma(0.5 * t2, t, (2.0 / 3.0) - t2);
    w2 = fma(0.5 * u2, um, (2.0 / 3.0) - u2);
}

__device__ __forceinline__ double4 ld4(const double4* p) {   // read-only path, two 128-bit loads
    const double2 a = __ldg(reinterpret_cast<const double2*>(p));
    const double2 b = __ldg(reinterpret_cast<const double2*>(p) + 1);
    return make_double4(a.x, a.y, b.x, b.y);
}

// u-side rows: one thread per l1 node.
__global__ void n0_uside_kernel(N0Params P, const double* __restrict__ l1, const double* __restrict__ w_l1,
                                const int* __restrict__ rowflag, const ClRow* __restrict__ rows, URow* __restrict__ utab) {
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= P.n_l1) return;
    double l = l1[r];
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (l >= 0.0 && l <= (double)(P.n_ell - 1)) {
        int i = (int)l;
        if (i > P.n_ell - 2) i = P.n_ell - 2;
        double w0, w1, w2, w3;
        bspline_weights(l - (double)i, w0, w1, w2, w3);
        const double* p = reinterpret_cast<const double*>(rows + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = w0 * p[k] + w1 * p[8 + k] + w2 * p[16 + k] + w3 * p[24 + k];
    }
    double e = exp(l * (l + 1.0) * P.a);
    URow u;
    u.l = l; u.wl = w_l1[r] * l;
    u.uT = v[0]; u.uX = v[1]; u.uE = v[2]; u.uB = v[3];
    u.T1 = v[4] + P.nT * e; u.X1 = v[5]; u.E1 = v[6] + P.nP * e;
    double B1 = v[7] + P.nP * e;
    u.rT1h = 0.5 / u.T1; u.rE1h = 0.5 / u.E1; u.rE1 = 1.0 / u.E1; u.rB1h = 0.5 / B1;
    u.flags = (double)rowflag[r];
    u.pad0 = u.pad1 = 0.0;
    utab[r] = u;
}

struct UVals { double uT, uX, uE, uB, T1, E1, X1, rT1h, rE1h, rE1, rB1h; };

// One integrand node (cmb/Nkk.py:66-162) for |u| = l at angle (c, s) = (cos phi, sin phi) and lensing multipole L.
// actT / actP: the |u| side of prep_lmaxmask (cmb/Nkk.py:17,20) for the 3000 / 5000 families; wl = weight * l.
template <bool SOA = false>
__device__ __forceinline__ void n0_node(const N0Params& P, const ClRow* __restrict__ rows, double L, double l, double c, double s,
                                        const UVals& u, bool actT, bool actP, double wl, bool mirror, double cm, double sm_, double fj,
                                        double& aTT, double& aTE, double& aEE, double& aEB, double& aBB) {
    // geometry exactly as cmb/Nkk.py:66-72,80 (no FMA contraction: |v| decides the masks)
    const double ux = __dmul_rn(l, c), uy = __dmul_rn(l, s);
    const double vx = __dsub_rn(L, ux);
    const double nv2 = __dadd_rn(__dmul_rn(vx, vx), __dmul_rn(uy, uy));
    const double nv = __dsqrt_rn(nv2);
    const bool vok = !(nv < P.lmin_v);
    bool okT = actT && vok && !(nv > P.lmax_T);
    bool okP = actP && vok && !(nv > P.lmax_P);
    double wT = wl, wP = wl;
    if (mirror) {
        // Folded phi grid: this node stands for (phi_j, 2pi - phi_j).  The two differ only by the rounding of cos/sin, which
        // matters when |v| sits on a cut (e.g. the 3000-4000-5000 triangle at phi = pi/2).  Integer test on the high word of
        // |v|^2 (ALU pipe, not FP64); inside the window both masks are evaluated exactly as the reference does.
        const unsigned hi = (unsigned)__double2hiint(nv2);
        if ((hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] || (hi - P.near_lo[2]) <= P.near_span[2]) {
            const double fm = 1.0 - fj;
            const double uxm = __dmul_rn(l, cm), uym = __dmul_rn(l, sm_);
            const double vxm = __dsub_rn(L, uxm);
            const double nvm = __dsqrt_rn(__dadd_rn(__dmul_rn(vxm, vxm), __dmul_rn(uym, uym)));
            const bool vokm = !(nvm < P.lmin_v);
            const bool okTm = actT && vokm && !(nvm > P.lmax_T);
            const bool okPm = actP && vokm && !(nvm > P.lmax_P);
            wT = wl * ((okT ? fj : 0.0) + (okTm ? fm : 0.0));
            wP = wl * ((okP ? fj : 0.0) + (okPm ? fm : 0.0));
            okT = okT || okTm;
            okP = okP || okPm;
        }
    }
    if (!(okT || okP)) return;

    const double Lx = ux + vx;                     // cmb/Nkk.py:82
    const double dLu = ux * Lx, dLv = vx * Lx;     // :84-85
    // cos 2phi_uv and sin^2 2phi_uv from dot / cross products instead of two atan2 (:74-78,106,130,138)
    const double dot = ux * vx - uy * uy, crs = uy * Lx;
    const double d2 = dot * dot, x2 = crs * crs;
    const double rden = 1.0 / (d2 + x2);
    const double c2 = (d2 - x2) * rden;
    const double s2sq = 4.0 * d2 * x2 * rden * rden;

    // eight cubic splines at |v| (prep_camb.py:88-89), zero outside the table
    int i = (int)nv;
    if (i > P.n_ell - 2) i = P.n_ell - 2;
    double w0, w1, w2, w3;
    bspline_weights(nv - (double)i, w0, w1, w2, w3);
    if (nv > (double)(P.n_ell - 1)) { w0 = w1 = w2 = w3 = 0.0; }
    double vT, vX, vE, vB, T2, X2, E2, B2;
    if (SOA) {
        const double* q = reinterpret_cast<const double*>(rows) + i;
        const size_t st = (size_t)P.soa_stride;
#define LBG_SPL(k) fma(w3, __ldg(q + (k) * st + 3), fma(w2, __ldg(q + (k) * st + 2), fma(w1, __ldg(q + (k) * st + 1), w0 * __ldg(q + (k) * st))))
        vT = LBG_SPL(0); vX = LBG_SPL(1); vE = LBG_SPL(2); vB = LBG_SPL(3);
        T2 = LBG_SPL(4); X2 = LBG_SPL(5); E2 = LBG_SPL(6); B2 = LBG_SPL(7);
#undef LBG_SPL
    } else {
        const double4* rp = reinterpret_cast<const double4*>(rows + i);
        double4 A = ld4(rp), B = ld4(rp + 1);
        vT = w0 * A.x; vX = w0 * A.y; vE = w0 * A.z; vB = w0 * A.w;
        T2 = w0 * B.x; X2 = w0 * B.y; E2 = w0 * B.z; B2 = w0 * B.w;
        A = ld4(rp + 2); B = ld4(rp + 3);
        vT = fma(w1, A.x, vT); vX = fma(w1, A.y, vX); vE = fma(w1, A.z, vE); vB = fma(w1, A.w, vB);
        T2 = fma(w1, B.x, T2); X2 = fma(w1, B.y, X2); E2 = fma(w1, B.z, E2); B2 = fma(w1, B.w, B2);
        A = ld4(rp + 4); B = ld4(rp + 5);
        vT = fma(w2, A.x, vT); vX = fma(w2, A.y, vX); vE = fma(w2, A.z, vE); vB = fma(w2, A.w, vB);
        T2 = fma(w2, B.x, T2); X2 = fma(w2, B.y, X2); E2 = fma(w2, B.z, E2); B2 = fma(w2, B.w, B2);
        A = ld4(rp + 6); B = ld4(rp + 7);
        vT = fma(w3, A.x, vT); vX = fma(w3, A.y, vX); vE = fma(w3, A.z, vE); vB = fma(w3, A.w, vB);
        T2 = fma(w3, B.x, T2); X2 = fma(w3, B.y, X2); E2 = fma(w3, B.z, E2); B2 = fma(w3, B.w, B2);
    }
    // detector noise at |v| (cmb/DetectorNoise.py:22-26)
    const double en = exp(nv * (nv + 1.0) * P.a);
    T2 = fma(P.nT, en, T2);
    E2 = fma(P.nP, en, E2);
    B2 = fma(P.nP, en, B2);

    const unsigned terms = P.terms;
    if (okT) {
        if (terms & LBG_TT) {                      // cmb/Nkk.py:93-99
            const double fa = u.uT * dLu + vT * dLv;
            aTT = fma(wT, fa * fa * u.rT1h / T2, aTT);
        }
        if (terms & LBG_TE) {                      // cmb/Nkk.py:101-125
            const double fa_uv = u.uX * c2 * dLu + vX * dLv;
            const double fa_vu = u.uX * dLu + vX * c2 * dLv;
            const double xx = u.X1 * X2;
            const double num = u.E1 * T2 * fa_uv - xx * fa_vu;
            const double den = u.T1 * E2 * u.E1 * T2 - xx * xx;
            aTE = fma(wT, fa_uv * num / den, aTE);
        }
    }
    if (okP) {
        if (terms & LBG_EE) {                      // cmb/Nkk.py:127-133
            const double fa = (u.uE * dLu + vE * dLv) * c2;
            aEE = fma(wP, fa * fa * u.rE1h / E2, aEE);
        }
        if (terms & LBG_EB) {                      // cmb/Nkk.py:135-146
            const double fa = u.uE * dLu - vB * dLv;
            aEB = fma(wP, fa * fa * s2sq * u.rE1 / B2, aEB);
        }
        if (terms & LBG_BB) {                      // cmb/Nkk.py:148-154
            const double fa = (u.uB * dLu + vB * dLv) * c2;
            aBB = fma(wP, fa * fa * u.rB1h / B2, aBB);
        }
    }
}

// Main kernel: regular rows (flag bit4 clear), phi nodes possibly folded onto mirror pairs.
__global__ void __launch_bounds__(128)
n0_kernel(N0Params P, const double* __restrict__ Ls, const double* __restrict__ phi /*[6][n_phi]: cos,sin,w,cos_m,sin_m,frac_j*/,
          const URow* __restrict__ utab, const ClRow* __restrict__ rows, double* __restrict__ partial) {
    const int npairs = P.n_L * P.n_phi;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < npairs;
    const int pp = valid ? p : npairs - 1;
    const int iL = pp / P.n_phi, j = pp - iL * P.n_phi;
    const double L = Ls[iL], c = phi[j], s = phi[P.n_phi + j];
    const int r0 = blockIdx.y * P.rows_per_chunk;
    const int r1 = min(P.n_l1, r0 + P.rows_per_chunk);
    const bool fold = P.fold != 0;
    const double cm = phi[3 * P.n_phi + j], sm_ = phi[4 * P.n_phi + j], fj = phi[5 * P.n_phi + j];
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;

    for (int r = r0; r < r1; ++r) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q3 = ld4(up + 3);                // rB1h, flags, nT(l), nP(l)
        const int fl = (int)q3.y;
        if (!(fl & 5) || (fl & 16)) continue;          // block-uniform: masked row, or handled by n0_special_kernel
        const double4 q0 = ld4(up);                    // l, wl, uT, uX
        const double4 q1 = ld4(up + 1);                // uE, uB, T1, E1
        const double4 q2 = ld4(up + 2);                // X1, rT1h, rE1h, rE1
        UVals u;
        u.uT = q0.z; u.uX = q0.w; u.uE = q1.x; u.uB = q1.y; u.T1 = q1.z; u.E1 = q1.w;
        u.X1 = q2.x; u.rT1h = q2.y; u.rE1h = q2.z; u.rE1 = q2.w; u.rB1h = q3.x;
        n0_node(P, rows, L, q0.x, c, s, u, fl & 1, fl & 4, q0.y, fold, cm, sm_, fj, aTT, aTE, aEE, aEB, aBB);
    }
    if (valid) {
        const double wphi = phi[2 * P.n_phi + j];
        const size_t stride = (size_t)P.n_chunk * npairs;
        const size_t o = (size_t)blockIdx.y * npairs + p;
        partial[o] = aTT * wphi;
        partial[stride + o] = aTE * wphi;
        partial[2 * stride + o] = aEE * wphi;
        partial[3 * stride + o] = aEB * wphi;
        partial[4 * stride + o] = aBB * wphi;
    }
}

// ---- "rows" variant: thread <-> one l1 row (consecutive lanes = consecutive |u|), loop over the phi nodes ------------
// The u-side values live in registers for the whole phi loop (no u-row traffic), cos/sin/weights of the phi nodes come
// from constant memory through the uniform datapath, and the four B-spline rows a warp gathers at |v| lie within a few
// cache lines because |v| moves by at most 0.5 between neighbouring lanes.
#define LBG_MAX_PHI_CONST 1024
__constant__ double c_phi[6 * LBG_MAX_PHI_CONST];   // [6][n_phi]: cos, sin, w, cos_m, sin_m, frac_j

#ifndef LBG_ROWS_MINB
#define LBG_ROWS_MINB 4
#endif
__global__ void __launch_bounds__(128, LBG_ROWS_MINB)
n0_rows_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const ClRow* __restrict__ rows,
               double* __restrict__ partial /*[5][n_L][n_rb]*/) {
    __shared__ double red[5][4];
    const int r = blockIdx.x * 128 + threadIdx.x;
    const int iL = blockIdx.y, n_rb = gridDim.x;
    const double L = Ls[iL];
    const bool fold = P.fold != 0;
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
    int fl = 0;
    double l = 1.0, wl = 0.0;
    UVals u = {};
    if (r < P.n_l1) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q0 = ld4(up), q1 = ld4(up + 1), q2 = ld4(up + 2), q3 = ld4(up + 3);
        fl = (int)q3.y;
        l = q0.x; wl = q0.y;
        u.uT = q0.z; u.uX = q0.w; u.uE = q1.x; u.uB = q1.y; u.T1 = q1.z; u.E1 = q1.w;
        u.X1 = q2.x; u.rT1h = q2.y; u.rE1h = q2.z; u.rE1 = q2.w; u.rB1h = q3.x;
    }
    if ((fl & 5) && !(fl & 16)) {
        const bool actT = fl & 1, actP = fl & 4;
        for (int j = 0; j < P.n_phi; ++j) {
            const double c = c_phi[j], s = c_phi[P.n_phi + j], wj = c_phi[2 * P.n_phi + j];
            const double cm = c_phi[3 * P.n_phi + j], sm_ = c_phi[4 * P.n_phi + j], fj = c_phi[5 * P.n_phi + j];
            n0_node<true>(P, rows, L, l, c, s, u, actT, actP, wj, fold, cm, sm_, fj, aTT, aTE, aEE, aEB, aBB);
        }
    }
    double acc[5] = {aTT * wl, aTE * wl, aEE * wl, aEB * wl, aBB * wl};
#pragma unroll
    for (int t = 0; t < 5; ++t) {
        double v = acc[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[t][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        const int t = threadIdx.x;
        const double v = (red[t][0] + red[t][1]) + (red[t][2] + red[t][3]);
        partial[((size_t)t * P.n_L + iL) * n_rb + blockIdx.x] = v;
    }
}

// One block per L: fixed-order sum of the row-block partials (+ the special rows), then cmb/Nkk.py:169-172.
__global__ void __launch_bounds__(128)
n0_rows_reduce_kernel(N0Params P, int n_rb, const double* __restrict__ Ls, const double* __restrict__ partial,
                      const double* __restrict__ spartial, int n_phi_full, double* __restrict__ out) {
    __shared__ double sm[5][128];
    const int iL = blockIdx.x;
    double acc[5] = {0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < n_rb; b += 128) {
#pragma unroll
        for (int t = 0; t < 5; ++t) acc[t] += partial[((size_t)t * P.n_L + iL) * n_rb + b];
    }
    if (spartial) {
        const size_t sstride = (size_t)P.n_L * n_phi_full;
        for (int j = threadIdx.x; j < n_phi_full; j += 128) {
#pragma unroll
            for (int t = 0; t < 5; ++t) acc[t] += spartial[t * sstride + (size_t)iL * n_phi_full + j];
        }
    }
#pragma unroll
    for (int t = 0; t < 5; ++t) sm[t][threadIdx.x] = acc[t];
    __syncthreads();
    for (int h = 64; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
#pragma unroll
            for (int t = 0; t < 5; ++t) sm[t][threadIdx.x] += sm[t][threadIdx.x + h];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double L = Ls[iL];
        double tot = (((sm[0][0] + sm[1][0]) + sm[2][0]) + P.ebfac * sm[3][0]) + sm[4][0];
        const double twopi = 2.0 * 3.14159265358979323846;
        double r = tot / (twopi * twopi);
        r = 1.0 / r;
        const double pre = L * (L + 1.0) / 2.0;
        out[iL] = r * (pre * pre);
        for (int t = 0; t < 5; ++t) out[(size_t)(t + 1) * P.n_L + iL] = sm[t][0];
    }
}

// ---- "rows2" variant: the rows decomposition with a restructured node ----------------------------------------------
// Same integrand, arranged for the FP64 pipe (the limiter once the table gathers are coalesced):
//  * |v|^2 = l (l - 2 L cos) + L^2 by two FMAs; one rsqrt yields both |v| and 1/|v|^2 (the cos 2phi_uv / sin^2 2phi_uv
//    denominators are |u|^2 |v|^2);  L.u, L.v, u.v, u x v come from per-(L, phi) products staged in shared memory;
//  * the four filter denominators (C^TT_tot(v), TE determinant, C^EE_tot(v), C^BB_tot(v)) share ONE reciprocal;
//  * detector noise at |v| = G[i] exp(a t (2 i + 1 + t)) with G[i] = exp(a i (i + 1)) tabulated per knot and a degree-7
//    Taylor polynomial in the small remainder (|x| < 0.03, checked on the host; otherwise exp() is called);
//  * masks are integer tests on the high word of |v|^2.  A node whose |v|^2 lies within 4e-6 of a cut is left to a second
//    pass that evaluates it with the bit-faithful n0_node (the reference decides those masks on its own rounding).
// Blocks whose rows all lie above lmax_T skip the temperature estimators and their four splines (DO_T = false).
struct PhiNode { double A, B, C, w; };     // 2 L cos phi, L cos phi, L sin phi, quadrature weight

__device__ __forceinline__ double rsqrt_fast(double x) {       // MUFU.RSQ64H seed + one third-order step: ~1 ulp, x normal > 0
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-(x * y), y, 1.0);
    const double p = fma(0.375, e, 0.5) * e;
    return fma(y, p, y);
}

__device__ __forceinline__ double rcp_fast(double x) {         // MUFU.RCP64H seed + one third-order step, x normal
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x, y, 1.0);
    const double q = fma(e, e, e);
    return fma(y, q, y);
}

template <bool DO_T, bool EXPN>
__device__ __forceinline__ void n0_fast_node(const N0Params& P, const double* __restrict__ tab, const double* __restrict__ G,
                                             double L2, double l, double l2, double rl2, const UVals& u, bool actT, bool actP,
                                             const PhiNode ph, double& aTT, double& aTE, double& aEE, double& aEB) {
    const double nv2 = fma(l, l - ph.A, L2);
    const unsigned hi = (unsigned)__double2hiint(nv2);
    const bool rare = (hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] || (hi - P.near_lo[2]) <= P.near_span[2];
    const bool vok = !rare && hi > P.near_lo[0];
    const bool okT = DO_T && actT && vok && hi < P.near_lo[1];
    const bool okP = actP && vok && hi < P.near_lo[2];
    if (!__any_sync(0xffffffffu, okT || okP)) return;

    const double y = rsqrt_fast(fmax(nv2, 1e-300));
    const double nv = nv2 * y;
    const double dLu = l * ph.B, dLv = L2 - dLu;           // L.u, L.v
    const double dot = dLu - l2, crs = l * ph.C;           // u.v, |u x v|
    const double d2 = dot * dot, x2 = crs * crs;
    const double rden = rl2 * (y * y);                     // 1 / (|u|^2 |v|^2)
    const double c2 = (d2 - x2) * rden;
    const double s2sq = fma(-c2, c2, 1.0);

    int i = min((int)nv, P.n_ell - 2);
    const double fi = (double)i;
    const double t = nv - fi, um = 1.0 - t;
    const double t2 = t * t, u2 = um * um;
    const double w0 = (u2 * (1.0 / 6.0)) * um, w3 = (t2 * (1.0 / 6.0)) * t;
    const double w1 = fma(t2, fma(0.5, t, -1.0), 2.0 / 3.0), w2 = fma(u2, fma(0.5, um, -1.0), 2.0 / 3.0);
    const double2* q = reinterpret_cast<const double2*>(tab) + i;   // (c_i, c_{i+1}) pairs, compile-time spectrum stride
    double vT = 0.0, vX = 0.0, T2 = 1.0, X2 = 0.0, vE, vB, E2, B2;
    {
        double2 a, b;
#define LBG_SPL(k, out) a = __ldg(q + (k) * LBG_PAIR_STRIDE); b = __ldg(q + (k) * LBG_PAIR_STRIDE + 2); \
                        out = fma(w3, b.y, fma(w2, b.x, fma(w1, a.y, w0 * a.x)))
        if (DO_T) { LBG_SPL(0, vT); LBG_SPL(1, vX); LBG_SPL(4, T2); LBG_SPL(5, X2); }
        LBG_SPL(2, vE); LBG_SPL(3, vB); LBG_SPL(6, E2); LBG_SPL(7, B2);
#undef LBG_SPL
    }
    double en;
    if (EXPN) {
        en = exp((nv2 + nv) * P.a);
    } else {
        const double x = (P.a * t) * fma(2.0, fi, 1.0 + t);
        double pl = fma(x, 1.0 / 5040.0, 1.0 / 720.0);
        pl = fma(pl, x, 1.0 / 120.0); pl = fma(pl, x, 1.0 / 24.0); pl = fma(pl, x, 1.0 / 6.0);
        pl = fma(pl, x, 0.5); pl = fma(pl, x, 1.0); pl = fma(pl, x, 1.0);
        en = __ldg(G + i) * pl;
    }
    if (DO_T) T2 = fma(P.nT, en, T2);
    E2 = fma(P.nP, en, E2);
    B2 = fma(P.nP, en, B2);

    // one reciprocal for the four filter denominators
    double E1T2 = 1.0, xx = 0.0, den = 1.0;
    if (DO_T) {
        E1T2 = u.E1 * T2; xx = u.X1 * X2;
        den = fma(u.T1 * E2, E1T2, -(xx * xx));
        if (!(P.terms & LBG_TE)) den = 1.0;
        if (!(P.terms & LBG_TT)) T2 = 1.0;
    }
    if (!(P.terms & LBG_EB)) B2 = 1.0;
    const double p1 = DO_T ? T2 * den : 1.0, p2 = E2 * B2;
    const double r = rcp_fast(DO_T ? p1 * p2 : p2);
    const double r34 = DO_T ? r * p1 : r;                  // 1 / (E2 B2)
    const double rE2 = r34 * B2, rB2 = r34 * E2;
    const double uEd = u.uE * dLu;
    if (DO_T) {
        const double r12 = r * p2;                         // 1 / (T2 den)
        const double rT2 = r12 * den, rdn = r12 * T2;
        const double fa = fma(vT, dLv, u.uT * dLu);
        const double tt = (fa * fa) * (u.rT1h * rT2);
        const double uXd = u.uX * dLu, vXd = vX * dLv;
        const double fa_uv = fma(uXd, c2, vXd), fa_vu = fma(vXd, c2, uXd);
        const double num = fma(E1T2, fa_uv, -(xx * fa_vu));
        const double te = (fa_uv * num) * rdn;
        aTT = fma(ph.w, okT ? tt : 0.0, aTT);
        aTE = fma(ph.w, okT ? te : 0.0, aTE);
    }
    const double fe = fma(vE, dLv, uEd) * c2;
    const double ee = (fe * fe) * (u.rE1h * rE2);
    const double fb = fma(-vB, dLv, uEd);
    const double eb = ((fb * fb) * s2sq) * (u.rE1 * rB2);
    aEE = fma(ph.w, okP ? ee : 0.0, aEE);
    aEB = fma(ph.w, okP ? eb : 0.0, aEB);
}

// G[i] = exp(a i (i + 1)) on the integer knots (cmb/DetectorNoise.py:22 without the amplitude).
// pairs[k][r] = (c_k[r], c_k[r + 1]) for the eight spectra (rows2 kernel): double2[8][LBG_PAIR_STRIDE], zero beyond the table.
__global__ void n0_pairs_kernel(int n, int stride, const double* __restrict__ soa, double* __restrict__ pairs) {
    const int r = blockIdx.x * blockDim.x + threadIdx.x, k = blockIdx.y;
    if (r >= n || r >= LBG_PAIR_STRIDE) return;
    pairs[((size_t)k * LBG_PAIR_STRIDE + r) * 2] = soa[(size_t)k * stride + r];
    pairs[((size_t)k * LBG_PAIR_STRIDE + r) * 2 + 1] = r + 1 < n ? soa[(size_t)k * stride + r + 1] : 0.0;
}

__global__ void n0_noise_knots_kernel(int n, double a, double* __restrict__ G) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) G[i] = exp(((double)i * ((double)i + 1.0)) * a);
}

template <bool DO_T, bool EXPN>
__device__ __forceinline__ void n0_rows2_loop(const N0Params& P, const double* __restrict__ tab, const double* __restrict__ G,
                                              const PhiNode* __restrict__ sph, double L2, double l, const UVals& u, bool actT, bool actP,
                                              double& aTT, double& aTE, double& aEE, double& aEB) {
    const double l2 = l * l, rl2 = 1.0 / l2;
    int j = 0;
    for (; j + 1 < P.n_phi; j += 2) {      // two independent nodes per trip: ILP for the FP64 pipe
        n0_fast_node<DO_T, EXPN>(P, tab, G, L2, l, l2, rl2, u, actT, actP, sph[j], aTT, aTE, aEE, aEB);
        n0_fast_node<DO_T, EXPN>(P, tab, G, L2, l, l2, rl2, u, actT, actP, sph[j + 1], aTT, aTE, aEE, aEB);
    }
    if (j < P.n_phi) n0_fast_node<DO_T, EXPN>(P, tab, G, L2, l, l2, rl2, u, actT, actP, sph[j], aTT, aTE, aEE, aEB);
}

template <bool EXPN>
__global__ void __launch_bounds__(128)
n0_rows2_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const double* __restrict__ tab /*pairs*/,
                const double* __restrict__ soa, const double* __restrict__ G, double* __restrict__ partial /*[5][n_L][n_rb]*/) {
    extern __shared__ double shm[];
    PhiNode* sph = reinterpret_cast<PhiNode*>(shm);                 // [n_phi]
    double (*red)[4] = reinterpret_cast<double (*)[4]>(shm + 4 * P.n_phi);
    __shared__ int s_anyT;
    const int r = blockIdx.x * 128 + threadIdx.x;
    const int iL = blockIdx.y, n_rb = gridDim.x;
    const double L = Ls[iL];
    const double L2 = L * L;
    if (threadIdx.x == 0) s_anyT = 0;
    for (int j = threadIdx.x; j < P.n_phi; j += 128) {
        const double c = c_phi[j], s = c_phi[P.n_phi + j];
        PhiNode n;
        n.B = L * c; n.A = 2.0 * n.B; n.C = L * s; n.w = c_phi[2 * P.n_phi + j];
        sph[j] = n;
    }
    int fl = 0;
    double l = 1.0, wl = 0.0;
    UVals u = {};
    if (r < P.n_l1) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q0 = ld4(up), q1 = ld4(up + 1), q2 = ld4(up + 2), q3 = ld4(up + 3);
        fl = (int)q3.y;
        l = q0.x; wl = q0.y;
        u.uT = q0.z; u.uX = q0.w; u.uE = q1.x; u.uB = q1.y; u.T1 = q1.z; u.E1 = q1.w;
        u.X1 = q2.x; u.rT1h = q2.y; u.rE1h = q2.z; u.rE1 = q2.w; u.rB1h = q3.x;
    }
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    __syncthreads();
    if (actT) s_anyT = 1;                                           // benign race: every writer stores 1
    __syncthreads();
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
    if (s_anyT) n0_rows2_loop<true, EXPN>(P, tab, G, sph, L2, l, u, actT, actP, aTT, aTE, aEE, aEB);
    else        n0_rows2_loop<false, EXPN>(P, tab, G, sph, L2, l, u, actT, actP, aTT, aTE, aEE, aEB);
    // second pass: nodes within 4e-6 of a cut in |v|^2, evaluated exactly as the reference rounds them
    if (act) {
        const bool fold = P.fold != 0;
        for (int j = 0; j < P.n_phi; ++j) {
            const double nv2 = fma(l, l - sph[j].A, L2);
            const unsigned hi = (unsigned)__double2hiint(nv2);
            if ((hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] || (hi - P.near_lo[2]) <= P.near_span[2])
                n0_node<true>(P, reinterpret_cast<const ClRow*>(soa), L, l, c_phi[j], c_phi[P.n_phi + j], u, fl & 1, fl & 4, sph[j].w, fold,
                              c_phi[3 * P.n_phi + j], c_phi[4 * P.n_phi + j], c_phi[5 * P.n_phi + j], aTT, aTE, aEE, aEB, aBB);
        }
    }
    if (!(P.terms & LBG_TT)) aTT = 0.0;
    if (!(P.terms & LBG_TE)) aTE = 0.0;
    if (!(P.terms & LBG_EE)) aEE = 0.0;
    if (!(P.terms & LBG_EB)) aEB = 0.0;
    double acc[5] = {aTT * wl, aTE * wl, aEE * wl, aEB * wl, aBB * wl};
#pragma unroll
    for (int t = 0; t < 5; ++t) {
        double v = acc[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[t][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        const int t = threadIdx.x;
        const double v = (red[t][0] + red[t][1]) + (red[t][2] + red[t][3]);
        partial[((size_t)t * P.n_L + iL) * n_rb + blockIdx.x] = v;
    }
}

// ---- "rows3" variant: rows decomposition, per-interval polynomial tables with the detector noise folded in ----------
// What bounds rows2 is not one pipe but three at once (ncu r1b: FP64 56 %, L1 data pipe 63 %, issue 56 %): every coefficient
// a lane loads costs register-fill bandwidth (128 B/clk/SM, no dedupe between lanes), and the B-spline form spends 12 FP64
// instructions on the basis weights plus 4 per spectrum.  rows3 therefore
//  * evaluates each spectrum from the monomial coefficients of its knot interval (Horner, 3 DFMA, one 32-byte row slot),
//    slot-major with a compile-time slot stride: the L1 data stage spends one wavefront per 128-byte line a load touches, so
//    neighbouring lanes (neighbouring |v|) must find their coefficients of one slot in neighbouring bytes;
//  * folds the detector noise into the C^tot slots when the table is built for the call: on [i, i+1] the factor
//    exp(a t (2i+1+t)) - 1 is replaced by its cubic interpolant at the four Chebyshev nodes (relative error bounded by
//    (beta^4 + 12 a beta^2 + 12 a^2) / 3072, beta = a (2 lmax + 1); the host falls back to rows2 when the bound exceeds 2e-12);
//  * drops the unlensed-BB slot when that spectrum is identically zero (CAMB scalar runs: prep_camb.py:84-85);
//  * takes cos phi_j / w_j from constant memory (uniform datapath) and needs only cos: cos 2phi_uv = 2 (u.v)^2/(|u|^2|v|^2) - 1;
//  * skips the temperature estimators (four slots, ~40 FP64 instructions) for warps whose |v| all exceed lmax_T.
// Nodes within 4e-6 of a cut in |v|^2 are still left to the bit-faithful second pass (n0_node).
__constant__ double2 c_cw[LBG_MAX_PHI_CONST];        // (cos phi_j, w_j) of the (folded) phi grid

struct ChebFit { double m[4][4]; double t[4]; };     // monomial coefficients = m . samples at t[]

// slots per interval: 0 C^EE_tot, 1 C^BB_tot, 2 unlensed EE, 3 unlensed BB | 4 C^TT_tot, 5 lensed TE, 6 unlensed TT, 7 unlensed TE
__global__ void n0_poly_table_kernel(N0Params P, ChebFit F, int n_int, const double* __restrict__ soa, double4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_int) return;
    const size_t st = (size_t)P.soa_stride;
    double4 a[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const double c0 = soa[k * st + i], c1 = soa[k * st + i + 1], c2 = soa[k * st + i + 2], c3 = soa[k * st + i + 3];
        a[k].x = (c0 + 4.0 * c1 + c2) * (1.0 / 6.0);
        a[k].y = (c2 - c0) * 0.5;
        a[k].z = (c0 - 2.0 * c1 + c2) * 0.5;
        a[k].w = ((c3 - c0) + 3.0 * (c1 - c2)) * (1.0 / 6.0);
    }
    const double fi = (double)i;
    double g[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) g[q] = expm1((P.a * F.t[q]) * fma(2.0, fi, 1.0 + F.t[q]));
    const double G = exp((fi * (fi + 1.0)) * P.a);
    double n[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) n[q] = G * (F.m[q][0] * g[0] + F.m[q][1] * g[1] + F.m[q][2] * g[2] + F.m[q][3] * g[3]);
    n[0] += G;
    // layout double[slot][coefficient][LBG_PAIR_STRIDE]: four 8-byte loads per slot.  scripts/micro/ldg_pattern.cu (B200): a warp
    // LDG.E.64 whose lane pairs share an element costs 2.6 SM-clk on the L1-hit path, an LDG.E.128 of the same shape 6.9 -- two 64-bit
    // loads move the same bytes 24 % cheaper, and the 16-17 intervals a warp touches are 136 contiguous bytes of one array
    double* o = reinterpret_cast<double*>(out) + i;
#define LBG_PUT(slot, v) o[(4 * (slot)) * LBG_PAIR_STRIDE] = (v).x; o[(4 * (slot) + 1) * LBG_PAIR_STRIDE] = (v).y; \
                         o[(4 * (slot) + 2) * LBG_PAIR_STRIDE] = (v).z; o[(4 * (slot) + 3) * LBG_PAIR_STRIDE] = (v).w
    const double4 e2 = make_double4(fma(P.nP, n[0], a[6].x), fma(P.nP, n[1], a[6].y), fma(P.nP, n[2], a[6].z), fma(P.nP, n[3], a[6].w));
    const double4 b2 = make_double4(fma(P.nP, n[0], a[7].x), fma(P.nP, n[1], a[7].y), fma(P.nP, n[2], a[7].z), fma(P.nP, n[3], a[7].w));
    const double4 t2 = make_double4(fma(P.nT, n[0], a[4].x), fma(P.nT, n[1], a[4].y), fma(P.nT, n[2], a[4].z), fma(P.nT, n[3], a[4].w));
    LBG_PUT(0, e2); LBG_PUT(1, b2); LBG_PUT(2, a[2]); LBG_PUT(3, a[3]);
    LBG_PUT(4, t2); LBG_PUT(5, a[5]); LBG_PUT(6, a[0]); LBG_PUT(7, a[1]);
#undef LBG_PUT
}

#ifndef LBG_ESTRIN
#define LBG_ESTRIN 0
#endif
template <int SLOT>
__device__ __forceinline__ double horner3(const double* __restrict__ p, double t, double t2) {    // p = table + interval index
    const double a0 = __ldg(p + (4 * SLOT) * LBG_PAIR_STRIDE), a1 = __ldg(p + (4 * SLOT + 1) * LBG_PAIR_STRIDE);
    const double a2 = __ldg(p + (4 * SLOT + 2) * LBG_PAIR_STRIDE), a3 = __ldg(p + (4 * SLOT + 3) * LBG_PAIR_STRIDE);
#if LBG_ESTRIN
    return fma(fma(a3, t, a2), t2, fma(a1, t, a0));               // two independent FMAs, then one: dependent depth 2 instead of 3
#else
    return fma(fma(fma(a3, t, a2), t, a1), t, a0);
#endif
}

struct RowGeom { double L2, l2, rl2x2, lL, m2lL, l2pL2; };   // per-thread constants of one (L, |u|) row

template <bool DO_T, bool HAS_UB, bool FULL>
__device__ __forceinline__ void n0_poly_node(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                             bool actT, bool actP, const double2 cw, bool& any_rare,
                                             double& aTT, double& aTE, double& aEE, double& aEB) {
    const double nv2 = fma(g.m2lL, cw.x, g.l2pL2);               // |v|^2 = l^2 + L^2 - 2 l L cos phi
    const unsigned hi = (unsigned)__double2hiint(nv2);
    const bool rare = (hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] || (hi - P.near_lo[2]) <= P.near_span[2];
    any_rare = any_rare || rare;                                 // the bit-faithful second pass only runs for rows that saw one
    const bool vok = !rare && hi > P.near_lo[0];
    const bool okT = DO_T && actT && vok && hi < P.near_lo[1];
    const bool okP = actP && vok && hi < P.near_lo[2];
    if (!__any_sync(0xffffffffu, okT || okP)) return;
    const bool doT = DO_T && __any_sync(0xffffffffu, okT);       // warp-uniform

    const double y = rsqrt_fast(nv2);                            // lanes with a bad |v|^2 carry NaN; their terms are selected away
    const double nv = nv2 * y;
    const double dLu = g.lL * cw.x, dLv = g.L2 - dLu;            // L.u, L.v
    const double dot = dLu - g.l2;                               // u.v
    const double c2 = fma((dot * dot) * (y * y), g.rl2x2, -1.0); // cos 2phi_uv
    const double s2sq = fma(-c2, c2, 1.0);                       // sin^2 2phi_uv

    const int i = max(min((int)nv, P.n_ell - 2), 0);
    const double t = nv - (double)i;
    const double t2 = t * t;
    const double* row = reinterpret_cast<const double*>(ptab) + i;
    const double E2 = horner3<0>(row, t, t2), vE = horner3<2>(row, t, t2);
    double B2 = horner3<1>(row, t, t2);
    const double vB = HAS_UB ? horner3<3>(row, t, t2) : 0.0;
    if (!FULL) { if (!(P.terms & LBG_EB)) B2 = 1.0; }
    const double uEd = u.uE * dLu;
    const double p2 = E2 * B2;
    double rE2, rB2;
    if (doT) {
        double T2 = horner3<4>(row, t, t2);
        const double X2 = horner3<5>(row, t, t2), vT = horner3<6>(row, t, t2), vX = horner3<7>(row, t, t2);
        const double E1T2 = u.E1 * T2, xx = u.X1 * X2;
        double den = fma(u.T1 * E2, E1T2, -(xx * xx));
        if (!FULL) { if (!(P.terms & LBG_TE)) den = 1.0;  if (!(P.terms & LBG_TT)) T2 = 1.0; }
        const double p1 = T2 * den;
        const double r = rcp_fast(p1 * p2);                      // one reciprocal for the four filter denominators
        const double r34 = r * p1, r12 = r * p2;
        rE2 = r34 * B2; rB2 = r34 * E2;
        const double rT2 = r12 * den, rdn = r12 * T2;
        const double fa = fma(vT, dLv, u.uT * dLu);
        const double tt = (fa * fa) * (u.rT1h * rT2);
        const double uXd = u.uX * dLu, vXd = vX * dLv;
        const double fa_uv = fma(uXd, c2, vXd), fa_vu = fma(vXd, c2, uXd);
        const double num = fma(E1T2, fa_uv, -(xx * fa_vu));
        const double te = (fa_uv * num) * rdn;
        aTT = fma(cw.y, okT ? tt : 0.0, aTT);
        aTE = fma(cw.y, okT ? te : 0.0, aTE);
    } else {
        const double r = rcp_fast(p2);
        rE2 = r * B2; rB2 = r * E2;
    }
    const double fe = fma(vE, dLv, uEd) * c2;
    const double ee = (fe * fe) * (u.rE1h * rE2);
    const double fb = HAS_UB ? fma(-vB, dLv, uEd) : uEd;
    const double eb = ((fb * fb) * s2sq) * (u.rE1 * rB2);
    aEE = fma(cw.y, okP ? ee : 0.0, aEE);
    aEB = fma(cw.y, okP ? eb : 0.0, aEB);
}

#ifndef LBG_ROWS3_UNROLL
#define LBG_ROWS3_UNROLL 2
#endif
template <bool DO_T, bool HAS_UB, bool FULL>
__device__ __forceinline__ void n0_rows3_loop(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                              bool actT, bool actP, bool& any_rare, double& aTT, double& aTE, double& aEE, double& aEB) {
    int j = 0;
#if LBG_ROWS3_UNROLL >= 3
    for (; j + 2 < P.n_phi; j += 3) {
        n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j], any_rare, aTT, aTE, aEE, aEB);
        n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j + 1], any_rare, aTT, aTE, aEE, aEB);
        n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j + 2], any_rare, aTT, aTE, aEE, aEB);
    }
#endif
#if LBG_ROWS3_UNROLL >= 2
    for (; j + 1 < P.n_phi; j += 2) {      // two independent nodes per trip: ILP for the FP64 pipe
        n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j], any_rare, aTT, aTE, aEE, aEB);
        n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j + 1], any_rare, aTT, aTE, aEE, aEB);
    }
#endif
    for (; j < P.n_phi; ++j) n0_poly_node<DO_T, HAS_UB, FULL>(P, ptab, g, u, actT, actP, c_cw[j], any_rare, aTT, aTE, aEE, aEB);
}

#ifndef LBG_ROWS3_MINB
#define LBG_ROWS3_MINB 5
#endif
template <bool HAS_UB, bool FULL>
__global__ void __launch_bounds__(128, LBG_ROWS3_MINB)
n0_rows3_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const double4* __restrict__ ptab,
                const double* __restrict__ soa, double* __restrict__ partial /*[5][n_L][n_rb]*/) {
    __shared__ double red[5][4];
    __shared__ int s_anyT;
    const int r = blockIdx.x * 128 + threadIdx.x;
    const int iL = blockIdx.y, n_rb = gridDim.x;
    const double L = Ls[iL];
    if (threadIdx.x == 0) s_anyT = 0;
    int fl = 0;
    double l = 1.0, wl = 0.0;
    UVals u = {};
    if (r < P.n_l1) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q0 = ld4(up), q1 = ld4(up + 1), q2 = ld4(up + 2), q3 = ld4(up + 3);
        fl = (int)q3.y;
        l = q0.x; wl = q0.y;
        u.uT = q0.z; u.uX = q0.w; u.uE = q1.x; u.uB = q1.y; u.T1 = q1.z; u.E1 = q1.w;
        u.X1 = q2.x; u.rT1h = q2.y; u.rE1h = q2.z; u.rE1 = q2.w; u.rB1h = q3.x;
    }
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    __syncthreads();
    if (actT) s_anyT = 1;                                           // benign race: every writer stores 1
    __syncthreads();
    RowGeom g;
    g.L2 = L * L; g.l2 = l * l; g.rl2x2 = 2.0 / g.l2; g.lL = l * L; g.m2lL = -2.0 * g.lL; g.l2pL2 = g.l2 + g.L2;
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
    bool any_rare = false;
    if (s_anyT) n0_rows3_loop<true, HAS_UB, FULL>(P, ptab, g, u, actT, actP, any_rare, aTT, aTE, aEE, aEB);
    else        n0_rows3_loop<false, HAS_UB, FULL>(P, ptab, g, u, actT, actP, any_rare, aTT, aTE, aEE, aEB);
    // second pass: nodes within 4e-6 of a cut in |v|^2, evaluated exactly as the reference rounds them
    if (act && any_rare) {
        const bool fold = P.fold != 0;
        for (int j = 0; j < P.n_phi; ++j) {
            const double nv2 = fma(g.m2lL, c_cw[j].x, g.l2pL2);
            const unsigned hi = (unsigned)__double2hiint(nv2);
            if ((hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] || (hi - P.near_lo[2]) <= P.near_span[2])
                n0_node<true>(P, reinterpret_cast<const ClRow*>(soa), L, l, c_phi[j], c_phi[P.n_phi + j], u, fl & 1, fl & 4, c_cw[j].y, fold,
                              c_phi[3 * P.n_phi + j], c_phi[4 * P.n_phi + j], c_phi[5 * P.n_phi + j], aTT, aTE, aEE, aEB, aBB);
        }
    }
    if (!FULL) {
        if (!(P.terms & LBG_TT)) aTT = 0.0;
        if (!(P.terms & LBG_TE)) aTE = 0.0;
        if (!(P.terms & LBG_EE)) aEE = 0.0;
        if (!(P.terms & LBG_EB)) aEB = 0.0;
    }
    double acc[5] = {aTT * wl, aTE * wl, aEE * wl, aEB * wl, aBB * wl};
#pragma unroll
    for (int t = 0; t < 5; ++t) {
        double v = acc[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[t][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        const int t = threadIdx.x;
        const double v = (red[t][0] + red[t][1]) + (red[t][2] + red[t][3]);
        partial[((size_t)t * P.n_L + iL) * n_rb + blockIdx.x] = v;
    }
}

// ---- "rows5": rows3 tables and arithmetic, masks turned into phi-index ranges -------------------------------------------
// On the folded grid cos phi_j decreases with j, so |v|^2 = l^2 + L^2 - 2 l L cos phi_j (and its high word `hi`, the value
// every mask test of rows3 looks at) is non-decreasing in j: each mask is a RANGE of j.  A thread finds its range ends once
// per row by bisection on `hi` (six searches of ~7 steps, 2 instructions per node amortised):
//     [j0s, j0e)  nodes inside the |v| = lmin window      -> second pass;   fast path needs j >= j0e
//     [jT , jTe)  nodes inside the |v| = lmax_T window    -> second pass;   T estimators need j < jT
//     [jP , jPe)  nodes inside the |v| = lmax_P window    -> second pass;   P estimators need j < jP
// and the warp walks j over [min j0e, max jT) with the 7-slot body, then up to max jP with the 3-slot body.  Gains over
// rows3 (ncu r1b: the kernel is issue-bound -- FP64 instructions hold the dispatch port for two cycles -- so every
// non-FP64 instruction removed counts): no per-node window arithmetic / votes / branches (about 25 of 99 instructions
// per warp-node), fully masked nodes are never visited, and the branch-free bodies let the scheduler overlap the table
// loads of two nodes.  The selected values are bit-identical to rows3 for every node it evaluates.
__device__ __forceinline__ int first_j_hi_ge(const RowGeom& g, int n_phi, unsigned x) {   // first j with hi(j) >= x (n_phi if none)
    int lo = 0, hi = n_phi;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const unsigned h = (unsigned)__double2hiint(fma(g.m2lL, c_cw[mid].x, g.l2pL2));
        if (h >= x) hi = mid; else lo = mid + 1;
    }
    return lo;
}

// First half of a node: everything the table loads depend on.  Computed one trip AHEAD of its use (software pipelining): the
// loads of a trip then issue at the top of the loop body and the next trip's rsqrt / index chain fills their latency.
struct NodePre { double y, t; int i; };
__device__ __forceinline__ NodePre n0_prep5(const N0Params& P, const RowGeom& g, double c) {
    NodePre q;
    const double nv2 = fma(g.m2lL, c, g.l2pL2);                  // |v|^2; lanes outside their own range may carry junk: selected away
    q.y = rsqrt_fast(nv2);
    const double nv = nv2 * q.y;
    q.i = max(min((int)nv, P.n_ell - 2), 0);
    q.t = nv - (double)q.i;
    return q;
}

template <bool DO_T, bool HAS_UB, bool FULL>
__device__ __forceinline__ void n0_node5(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                         bool okT, bool okP, const double2 cw, const NodePre pre,
                                         double& aTT, double& aTE, double& aEE, double& aEB) {
    const double y = pre.y, t = pre.t;
    const double* row = reinterpret_cast<const double*>(ptab) + pre.i;
    const double dLu = g.lL * cw.x, dLv = g.L2 - dLu;            // L.u, L.v
    const double dot = dLu - g.l2;                               // u.v
    const double c2 = fma((dot * dot) * (y * y), g.rl2x2, -1.0); // cos 2phi_uv = 2 (u.v)^2 / (|u|^2 |v|^2) - 1
    const double s2sq = fma(-c2, c2, 1.0);                       // sin^2 2phi_uv
    const double E2 = horner3<0>(row, t, 0.0), vE = horner3<2>(row, t, 0.0);
    double B2 = horner3<1>(row, t, 0.0);
    const double vB = HAS_UB ? horner3<3>(row, t, 0.0) : 0.0;
    if (!FULL) { if (!(P.terms & LBG_EB)) B2 = 1.0; }
    const double uEd = u.uE * dLu;
    const double p2 = E2 * B2;
    double rE2, rB2;
    if (DO_T) {
        double T2 = horner3<4>(row, t, 0.0);
        const double X2 = horner3<5>(row, t, 0.0), vT = horner3<6>(row, t, 0.0), vX = horner3<7>(row, t, 0.0);
        const double E1T2 = u.E1 * T2, xx = u.X1 * X2;
        double den = fma(u.T1 * E2, E1T2, -(xx * xx));
        if (!FULL) { if (!(P.terms & LBG_TE)) den = 1.0;  if (!(P.terms & LBG_TT)) T2 = 1.0; }
        const double p1 = T2 * den;
        const double r = rcp_fast(p1 * p2);                      // one reciprocal for the four filter denominators
        const double r34 = r * p1, r12 = r * p2;
        rE2 = r34 * B2; rB2 = r34 * E2;
        const double rT2 = r12 * den, rdn = r12 * T2;
        const double fa = fma(vT, dLv, u.uT * dLu);
        const double tt = (fa * fa) * (u.rT1h * rT2);
        const double uXd = u.uX * dLu, vXd = vX * dLv;
        const double fa_uv = fma(uXd, c2, vXd), fa_vu = fma(vXd, c2, uXd);
        const double num = fma(E1T2, fa_uv, -(xx * fa_vu));
        const double te = (fa_uv * num) * rdn;
        aTT = fma(cw.y, okT ? tt : 0.0, aTT);
        aTE = fma(cw.y, okT ? te : 0.0, aTE);
    } else {
        const double r = rcp_fast(p2);
        rE2 = r * B2; rB2 = r * E2;
    }
    const double fe = fma(vE, dLv, uEd) * c2;
    const double ee = (fe * fe) * (u.rE1h * rE2);
    const double fb = HAS_UB ? fma(-vB, dLv, uEd) : uEd;
    const double eb = ((fb * fb) * s2sq) * (u.rE1 * rB2);
    aEE = fma(cw.y, okP ? ee : 0.0, aEE);
    aEB = fma(cw.y, okP ? eb : 0.0, aEB);
}

#ifndef LBG_ROWS5_MINB
#define LBG_ROWS5_MINB 4
#endif
#ifndef LBG_ROWS5_UNROLL
#define LBG_ROWS5_UNROLL 2
#endif
template <bool HAS_UB, bool FULL>
__global__ void __launch_bounds__(128, LBG_ROWS5_MINB)
n0_rows5_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const double4* __restrict__ ptab,
                const double* __restrict__ soa, double* __restrict__ partial /*[5][n_L][n_rb]*/) {
    __shared__ double red[5][4];
    const int r = blockIdx.x * 128 + threadIdx.x;
    const int iL = blockIdx.y, n_rb = gridDim.x;
    const double L = Ls[iL];
    int fl = 0;
    double l = 1.0, wl = 0.0;
    UVals u = {};
    if (r < P.n_l1) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q0 = ld4(up), q1 = ld4(up + 1), q2 = ld4(up + 2), q3 = ld4(up + 3);
        fl = (int)q3.y;
        l = q0.x; wl = q0.y;
        u.uT = q0.z; u.uX = q0.w; u.uE = q1.x; u.uB = q1.y; u.T1 = q1.z; u.E1 = q1.w;
        u.X1 = q2.x; u.rT1h = q2.y; u.rE1h = q2.z; u.rE1 = q2.w; u.rB1h = q3.x;
    }
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    RowGeom g;
    g.L2 = L * L; g.l2 = l * l; g.rl2x2 = 2.0 / g.l2; g.lL = l * L; g.m2lL = -2.0 * g.lL; g.l2pL2 = g.l2 + g.L2;
    const int n = P.n_phi;
    // range ends of this row (see the header); inactive lanes get empty ranges that do not widen the warp's walk
    int j0s = n, j0e = n, jT = 0, jTe = 0, jP = 0, jPe = 0;
    if (act) {
        j0s = first_j_hi_ge(g, n, P.near_lo[0]);
        j0e = first_j_hi_ge(g, n, P.near_lo[0] + P.near_span[0] + 1u);
        jT = first_j_hi_ge(g, n, P.near_lo[1]);
        jTe = first_j_hi_ge(g, n, P.near_lo[1] + P.near_span[1] + 1u);
        jP = first_j_hi_ge(g, n, P.near_lo[2]);
        jPe = first_j_hi_ge(g, n, P.near_lo[2] + P.near_span[2] + 1u);
    }
    const bool any_rare = act && (j0e > j0s || jTe > jT || jPe > jP);
    const int endT = actT ? jT : 0, endP = actP ? jP : 0;       // first j at which the T / P estimators stop (|v| side)
    const int w0 = __reduce_min_sync(0xffffffffu, j0e);
    const int wT = __reduce_max_sync(0xffffffffu, endT), wP = __reduce_max_sync(0xffffffffu, endP);
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
#define LBG_OKT(j) ((j) >= j0e && (j) < endT)
#define LBG_OKP(j) ((j) >= j0e && (j) < endP && !((j) >= jT && (j) < jTe))
    // pairs (j, j+1) from w0 on: 7-slot body while j < wT, then 3-slot body while j < wP.  A pair's second node may lie past a
    // range end (or past n_phi): its predicates are false there, and c_cw holds LBG_MAX_PHI_CONST finite entries.
    int j = w0;
    NodePre p0 = n0_prep5(P, g, c_cw[j].x), p1 = n0_prep5(P, g, c_cw[j + 1].x);
    for (; j < wT; j += 2) {
        const NodePre n0 = n0_prep5(P, g, c_cw[j + 2].x), n1 = n0_prep5(P, g, c_cw[j + 3].x);
        n0_node5<true, HAS_UB, FULL>(P, ptab, g, u, LBG_OKT(j), LBG_OKP(j), c_cw[j], p0, aTT, aTE, aEE, aEB);
        n0_node5<true, HAS_UB, FULL>(P, ptab, g, u, LBG_OKT(j + 1), LBG_OKP(j + 1), c_cw[j + 1], p1, aTT, aTE, aEE, aEB);
        p0 = n0; p1 = n1;
    }
    for (; j < wP; j += 2) {
        const NodePre n0 = n0_prep5(P, g, c_cw[j + 2].x), n1 = n0_prep5(P, g, c_cw[j + 3].x);
        n0_node5<false, HAS_UB, FULL>(P, ptab, g, u, false, LBG_OKP(j), c_cw[j], p0, aTT, aTE, aEE, aEB);
        n0_node5<false, HAS_UB, FULL>(P, ptab, g, u, false, LBG_OKP(j + 1), c_cw[j + 1], p1, aTT, aTE, aEE, aEB);
        p0 = n0; p1 = n1;
    }
#undef LBG_OKT
#undef LBG_OKP
    if (any_rare) {                                              // bit-faithful second pass over the window nodes only
        const bool fold = P.fold != 0;
        for (int k = 0; k < 3; ++k) {
            const int a = k == 0 ? j0s : (k == 1 ? jT : jP), b = k == 0 ? j0e : (k == 1 ? jTe : jPe);
            for (int jj = a; jj < b; ++jj) {
                if (k == 2 && jj >= jT && jj < jTe) continue;    // lmax_T == lmax_P: the windows coincide, visit once
                if (k == 1 && jj >= j0s && jj < j0e) continue;
                if (k == 2 && jj >= j0s && jj < j0e) continue;
                n0_node<true>(P, reinterpret_cast<const ClRow*>(soa), L, l, c_phi[jj], c_phi[n + jj], u, fl & 1, fl & 4, c_cw[jj].y, fold,
                              c_phi[3 * n + jj], c_phi[4 * n + jj], c_phi[5 * n + jj], aTT, aTE, aEE, aEB, aBB);
            }
        }
    }
    if (!FULL) {
        if (!(P.terms & LBG_TT)) aTT = 0.0;
        if (!(P.terms & LBG_TE)) aTE = 0.0;
        if (!(P.terms & LBG_EE)) aEE = 0.0;
        if (!(P.terms & LBG_EB)) aEB = 0.0;
    }
    double acc[5] = {aTT * wl, aTE * wl, aEE * wl, aEB * wl, aBB * wl};
#pragma unroll
    for (int t = 0; t < 5; ++t) {
        double v = acc[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[t][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 5) {
        const int t = threadIdx.x;
        const double v = (red[t][0] + red[t][1]) + (red[t][2] + red[t][3]);
        partial[((size_t)t * P.n_L + iL) * n_rb + blockIdx.x] = v;
    }
}

// Special rows, always on the full (unfolded) phi grid:
//  * rows sitting on the lmax_T / lmax_P cuts, where the reference's mask `|u| > lmax` (cmb/Nkk.py:17,20) depends on the
//    rounding of |u| = sqrt((l cos)^2 + (l sin)^2) at each phi node;
//  * rows the caller wants evaluated with its own u-side spline values at |u|_j (rows below the first non-zero C_l,
//    where the reference's filter 1/C^tot is pure noise and its result inherits the rounding of scipy's spline).
// su: [n_special][n_phi][8] splines at |u|_j (uTT,uTE,uEE,uBB,lTT,lTE,lEE,lBB); sm: [n_special][n_phi] |u|_j.
__global__ void __launch_bounds__(128)
n0_special_kernel(N0Params P, int n_special, int n_phi_full, const int* __restrict__ srow, const double* __restrict__ Ls,
                  const double* __restrict__ phi /*[3][n_phi_full]*/, const URow* __restrict__ utab,
                  const double* __restrict__ su, const double* __restrict__ snu, const ClRow* __restrict__ rows,
                  double* __restrict__ partial /*[5][n_L*n_phi_full]*/) {
    const int npairs = P.n_L * n_phi_full;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const int iL = p / n_phi_full, j = p - iL * n_phi_full;
    const double L = Ls[iL], c = phi[j], s = phi[n_phi_full + j];
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
    for (int k = 0; k < n_special; ++k) {
        const URow& ur = utab[srow[k]];
        const double* v = su + ((size_t)k * n_phi_full + j) * 8;
        const double nu = snu[(size_t)k * n_phi_full + j];
        const double en = exp(nu * (nu + 1.0) * P.a);
        UVals u;
        u.uT = v[0]; u.uX = v[1]; u.uE = v[2]; u.uB = v[3];
        u.T1 = v[4] + P.nT * en; u.X1 = v[5]; u.E1 = v[6] + P.nP * en;
        const double B1 = v[7] + P.nP * en;
        u.rT1h = 0.5 / u.T1; u.rE1h = 0.5 / u.E1; u.rE1 = 1.0 / u.E1; u.rB1h = 0.5 / B1;
        n0_node(P, rows, L, ur.l, c, s, u, !(nu > P.lmax_T), !(nu > P.lmax_P), ur.wl, false, 0.0, 0.0, 1.0, aTT, aTE, aEE, aEB, aBB);
    }
    const double wphi = phi[2 * n_phi_full + j];
    const size_t stride = (size_t)npairs;
    partial[p] = aTT * wphi;
    partial[stride + p] = aTE * wphi;
    partial[2 * stride + p] = aEE * wphi;
    partial[3 * stride + p] = aEB * wphi;
    partial[4 * stride + p] = aBB * wphi;
}

// Library-side u values for special rows without caller-supplied ones: the eight splines at |u|_j.
__global__ void n0_special_u_kernel(int n_ell, int n, const double* __restrict__ nu, const ClRow* __restrict__ rows,
                                    double* __restrict__ out /*[n][8]*/) {
    int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n) return;
    double x = nu[t];
    double v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (x >= 0.0 && x <= (double)(n_ell - 1)) {
        int i = (int)x;
        if (i > n_ell - 2) i = n_ell - 2;
        double w0, w1, w2, w3;
        bspline_weights(x - (double)i, w0, w1, w2, w3);
        const double* p = reinterpret_cast<const double*>(rows + i);
#pragma unroll
        for (int k = 0; k < 8; ++k) v[k] = w0 * p[k] + w1 * p[8 + k] + w2 * p[16 + k] + w3 * p[24 + k];
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) out[(size_t)t * 8 + k] = v[k];
}

// One block per L: fixed-order reduction of the partial sums, then cmb/Nkk.py:169-172.
__global__ void __launch_bounds__(128)
n0_reduce_kernel(N0Params P, const double* __restrict__ Ls, const double* __restrict__ partial,
                 const double* __restrict__ spartial, int n_phi_full,
                 double* __restrict__ out /*[6][n_L]: N0, TT, TE, EE, EB, BB integrals*/) {
    __shared__ double sm[5][128];
    const int iL = blockIdx.x, npairs = P.n_L * P.n_phi;
    const size_t stride = (size_t)P.n_chunk * npairs;
    const int n = P.n_chunk * P.n_phi;
    double acc[5] = {0, 0, 0, 0, 0};
    for (int e = threadIdx.x; e < n; e += 128) {
        int ch = e / P.n_phi, j = e - ch * P.n_phi;
        size_t o = (size_t)ch * npairs + (size_t)iL * P.n_phi + j;
#pragma unroll
        for (int t = 0; t < 5; ++t) acc[t] += partial[t * stride + o];
    }
    if (spartial) {
        const size_t sstride = (size_t)P.n_L * n_phi_full;
        for (int j = threadIdx.x; j < n_phi_full; j += 128) {
#pragma unroll
            for (int t = 0; t < 5; ++t) acc[t] += spartial[t * sstride + (size_t)iL * n_phi_full + j];
        }
    }
#pragma unroll
    for (int t = 0; t < 5; ++t) sm[t][threadIdx.x] = acc[t];
    __syncthreads();
    for (int h = 64; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
#pragma unroll
            for (int t = 0; t < 5; ++t) sm[t][threadIdx.x] += sm[t][threadIdx.x + h];
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        const double L = Ls[iL];
        double tot = (((sm[0][0] + sm[1][0]) + sm[2][0]) + P.ebfac * sm[3][0]) + sm[4][0];
        const double twopi = 2.0 * 3.14159265358979323846;
        double r = tot / (twopi * twopi);
        r = 1.0 / r;
        const double pre = L * (L + 1.0) / 2.0;
        out[iL] = r * (pre * pre);
        for (int t = 0; t < 5; ++t) out[(size_t)(t + 1) * P.n_L + iL] = sm[t][0];
    }
}
}  // namespace lbg

using namespace lbg;

extern "C" int lbg_n0_ho02(lbg_ctx* c, int n_L, const double* L, unsigned terms_mask,
                           double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                           int n_l1, const double* l1, const double* w_l1,
                           int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                           double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                           int n_exact, const int* exact_rows, const double* exact_u,
                           double* out_nkk, double* out_terms, long long* nodes_executed) {
    LBG_REQUIRE(c && L && l1 && w_l1 && cos_phi && sin_phi && w_phi && out_nkk, "lbg_n0_ho02: NULL argument");
    if (c->n_ell <= 0) { set_error("lbg_n0_ho02: call lbg_upload_cl_tables first"); return LBG_ESTATE; }
    LBG_REQUIRE(n_L >= 0 && n_l1 > 0 && n_phi > 0, "lbg_n0_ho02: bad sizes n_L=%d n_l1=%d n_phi=%d", n_L, n_l1, n_phi);
    LBG_REQUIRE((terms_mask & ~(LBG_TT | LBG_TE | LBG_EE | LBG_EB | LBG_BB)) == 0 && terms_mask != 0,
                "lbg_n0_ho02: unsupported terms mask 0x%x", terms_mask);
    LBG_REQUIRE(Tcmb_K > 0 && lmax_T <= lmax_P, "lbg_n0_ho02: need Tcmb > 0 and lmax_T <= lmax_P");
    LBG_REQUIRE(n_exact >= 0 && (n_exact == 0 || (exact_rows && exact_u)), "lbg_n0_ho02: exact rows given without values");
    if (n_L == 0) { if (nodes_executed) *nodes_executed = 0; return LBG_OK; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));

    // ---- phi nodes of the main kernel (optionally folded onto one node per mirror pair) -------------------
    std::vector<double> ph;   // [6][n_phe]: cos, sin, w (pair sum), mirror cos, mirror sin, w_j / (w_j + w_m)
    int n_phe = 0;
    if (fold_phi) {
        n_phe = (n_phi + 1) / 2;
        ph.assign(6 * (size_t)n_phe, 0.0);
        for (int j = 0, m = n_phi - 1; j <= m; ++j, --m) {
            bool ok = fabs(cos_phi[j] - cos_phi[m]) < 1e-9 && fabs(sin_phi[j] + sin_phi[m]) < 1e-9;
            LBG_REQUIRE(ok, "lbg_n0_ho02: fold_phi requested but nodes %d and %d are not mirror images", j, m);
            const double w = (m != j) ? w_phi[j] + w_phi[m] : w_phi[j];
            ph[j] = cos_phi[j]; ph[n_phe + j] = sin_phi[j]; ph[2 * (size_t)n_phe + j] = w;
            ph[3 * (size_t)n_phe + j] = cos_phi[m]; ph[4 * (size_t)n_phe + j] = sin_phi[m];
            ph[5 * (size_t)n_phe + j] = (m != j && w != 0.0) ? w_phi[j] / w : 1.0;
        }
    } else {
        n_phe = n_phi;
        ph.assign(6 * (size_t)n_phe, 0.0);
        memcpy(&ph[0], cos_phi, (size_t)n_phi * 8);
        memcpy(&ph[n_phe], sin_phi, (size_t)n_phi * 8);
        memcpy(&ph[2 * (size_t)n_phe], w_phi, (size_t)n_phi * 8);
    }
    std::vector<double> phf(3 * (size_t)n_phi);  // full grid for the special rows
    memcpy(&phf[0], cos_phi, (size_t)n_phi * 8);
    memcpy(&phf[n_phi], sin_phi, (size_t)n_phi * 8);
    memcpy(&phf[2 * (size_t)n_phi], w_phi, (size_t)n_phi * 8);

    // ---- row flags; special rows = rows on a cut of prep_lmaxmask + rows the caller supplies exact u values for
    std::vector<int> rowflag(n_l1, 0);
    std::vector<int> srow;                       // special rows, caller's first (same order as exact_u)
    for (int k = 0; k < n_exact; ++k) {
        LBG_REQUIRE(exact_rows[k] >= 0 && exact_rows[k] < n_l1, "lbg_n0_ho02: exact row %d out of range", exact_rows[k]);
        srow.push_back(exact_rows[k]);
    }
    long long active_rows = 0;
    for (int r = 0; r < n_l1; ++r) {
        double l = l1[r];
        LBG_REQUIRE(l > 0.0, "lbg_n0_ho02: l1 nodes must be positive");
        int f = 0;
        bool onT = fabs(l - lmax_T) <= 1e-12 * lmax_T, onP = fabs(l - lmax_P) <= 1e-12 * lmax_P;
        if (l < lmax_T || onT) f |= 1;
        if (l < lmax_P || onP) f |= 4;
        // a row sitting on the LAST KNOT of a table that ends inside the integration range: the reference's interp1d returns the
        // knot value or its fill value 0 per phi node, depending on whether |u|_j = sqrt((l cos)^2 + (l sin)^2) rounds to the knot
        // or one ulp above it (prep_camb.py:88-89) -- the same per-node rounding as the lmax cuts, so the same treatment
        const double l_end = (double)(c->n_ell - 1);
        const bool onEnd = f != 0 && fabs(l - l_end) <= 1e-12 * l_end;
        if (onT || onP || onEnd) {
            bool have = false;
            for (int q : srow) have = have || q == r;
            if (!have) srow.push_back(r);
        }
        rowflag[r] = f;
        if (f) ++active_rows;
    }
    for (int q : srow) rowflag[q] |= 16;
    const int n_special = (int)srow.size();
    // |u|_j exactly as cmb/Nkk.py:66-67,79 for every (special row, phi node)
    std::vector<double> snu((size_t)n_special * n_phi);
    for (int k = 0; k < n_special; ++k)
        for (int j = 0; j < n_phi; ++j) {
            volatile double ux = l1[srow[k]] * cos_phi[j], uy = l1[srow[k]] * sin_phi[j];
            volatile double a2 = ux * ux, b2 = uy * uy;
            volatile double s2 = a2 + b2;
            snu[(size_t)k * n_phi + j] = sqrt(s2);
        }

    // ---- noise constants, cmb/DetectorNoise.py:13-26 ------------------------------------------------------
    const double arcmin2rad = M_PI / (60. * 180.);
    double dT = deltaT_uK_arcmin * pow(10., -6.);
    dT *= arcmin2rad;
    const double tb = thetab_arcmin * arcmin2rad;
    const double dP = sqrt(2.) * dT;

    N0Params P;
    P.soa_stride = c->cl_soa_stride;
    P.n_L = n_L; P.n_phi = n_phe; P.n_l1 = n_l1; P.n_ell = c->n_ell; P.terms = terms_mask;
    P.nT = (dT / Tcmb_K) * (dT / Tcmb_K);
    P.nP = (dP / Tcmb_K) * (dP / Tcmb_K);
    P.a = tb * tb / (8. * log(2.));
    P.lmax_T = lmax_T; P.lmax_P = lmax_P; P.lmin_v = lmin_v;
    P.ebfac = iterative ? 2.5 : 1.0;                              // cmb/Nkk.py:141-143
    P.fold = fold_phi ? 1 : 0;
    {
        const double thr[3] = {lmin_v, lmax_T, lmax_P};
        for (int k = 0; k < 3; ++k) {
            double lo = thr[k] * thr[k] * (1.0 - 4e-6), hi = thr[k] * thr[k] * (1.0 + 4e-6);
            unsigned long long blo, bhi;
            memcpy(&blo, &lo, 8); memcpy(&bhi, &hi, 8);
            P.near_lo[k] = (unsigned)(blo >> 32);
            P.near_span[k] = (unsigned)(bhi >> 32) - P.near_lo[k];
        }
    }
    const long long npairs = (long long)n_L * n_phe, npairs_full = (long long)n_L * n_phi;
    LBG_REQUIRE(npairs_full < (1LL << 30), "lbg_n0_ho02: n_L * n_phi too large");
    const int bx = (int)((npairs + 127) / 128);
    int n_chunk = (c->sm_count * 32 + bx - 1) / bx;
    n_chunk = std::max(1, std::min(n_chunk, std::max(1, n_l1 / 128)));
    P.rows_per_chunk = (n_l1 + n_chunk - 1) / n_chunk;
    n_chunk = (n_l1 + P.rows_per_chunk - 1) / P.rows_per_chunk;
    P.n_chunk = n_chunk;

    // kernel variant: 0 = thread per (L, phi) marching over |u| rows, 1 = thread per |u| row looping over phi ("rows")
    int variant = 5;
    if (const char* e = getenv("LBG_N0_VARIANT")) variant = atoi(e);
    if (n_phe > LBG_MAX_PHI_CONST || n_L > 65535) variant = 0;
    // rows2 needs: only TT/TE/EE/EB requested, every |v| up to lmax_P inside the table, cuts ordered lmin_v < lmax_T <= lmax_P
    if (variant >= 2 && ((terms_mask & LBG_BB) || lmax_P > (double)(c->n_ell - 2) || c->n_ell + 2 > LBG_PAIR_STRIDE || !(lmin_v > 0.0 && lmin_v < lmax_T))) variant = 1;
    // rows5 walks index ranges: needs cos phi_j strictly decreasing (the folded grid) and disjoint, ordered cut windows
    bool rows5 = variant == 5;
    if (rows5) {
        variant = 3;
        for (int j = 1; j < n_phe; ++j) rows5 = rows5 && ph[j] < ph[j - 1];
        rows5 = rows5 && n_phe + 4 <= LBG_MAX_PHI_CONST && fold_phi && P.near_lo[0] + P.near_span[0] < P.near_lo[1] && P.near_lo[1] <= P.near_lo[2];
    }
    ChebFit fit;
    if (variant == 3) {
        // rows3 folds the noise into the table: relative error bound of the per-interval Chebyshev cubic (see the kernel's header)
        const double beta = P.a * (2.0 * lmax_P + 1.0);
        const double bound = (beta * beta * beta * beta + 12.0 * P.a * beta * beta + 12.0 * P.a * P.a) / 3072.0;
        if (bound > 2e-12) variant = 2;
        // monomial <- samples matrix of the four Chebyshev nodes on [0, 1] (Gauss-Jordan in long double)
        long double V[4][8];
        for (int q = 0; q < 4; ++q) {
            const long double tq = 0.5L + 0.5L * cosl((2 * q + 1) * M_PIl / 8.0L);
            fit.t[q] = (double)tq;
        }
        for (int q = 0; q < 4; ++q) {
            long double pw = 1.0L;
            for (int k = 0; k < 4; ++k) { V[q][k] = pw; pw *= (long double)fit.t[q]; V[q][4 + k] = (q == k) ? 1.0L : 0.0L; }
        }
        for (int col = 0; col < 4; ++col) {
            int piv = col;
            for (int q = col + 1; q < 4; ++q) if (fabsl(V[q][col]) > fabsl(V[piv][col])) piv = q;
            for (int k = 0; k < 8; ++k) std::swap(V[col][k], V[piv][k]);
            const long double d = V[col][col];
            for (int k = 0; k < 8; ++k) V[col][k] /= d;
            for (int q = 0; q < 4; ++q) if (q != col) { const long double f = V[q][col]; for (int k = 0; k < 8; ++k) V[q][k] -= f * V[col][k]; }
        }
        for (int q = 0; q < 4; ++q) for (int k = 0; k < 4; ++k) fit.m[q][k] = (double)V[q][4 + k];
    }
    const int n_rb = (n_l1 + 127) / 128;
    const size_t su_bytes = (size_t)n_special * n_phi * 8 * 8, snu_bytes = (size_t)n_special * n_phi * 8;
    if (c->n0_L.reserve((size_t)n_L * 8) || c->s0.reserve((size_t)n_l1 * 8) || c->s1.reserve((size_t)n_l1 * 8) ||
        c->n0_rowflag.reserve((size_t)n_l1 * 4) || c->n0_G.reserve((size_t)c->n_ell * 8 + 64) || c->n0_phi.reserve((ph.size() + phf.size()) * 8) ||
        c->n0_utab.reserve((size_t)n_l1 * sizeof(URow)) ||
        c->n0_partial.reserve((size_t)5 * (std::max((long long)n_chunk * npairs, (long long)n_L * n_rb) + npairs_full) * 8) || c->n0_out.reserve((size_t)6 * n_L * 8) ||
        (variant == 3 && c->cl_poly.reserve((size_t)LBG_PAIR_STRIDE * 8 * 32)) || c->s2.reserve(su_bytes + 64) || c->s3.reserve(snu_bytes + 64) || c->s4.reserve((size_t)n_special * 4 + 64))
        return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    double* d_ph = c->n0_phi.as<double>();
    double* d_phf = d_ph + ph.size();
    double* d_part = c->n0_partial.as<double>();
    double* d_spart = d_part + (size_t)5 * std::max((long long)n_chunk * npairs, (long long)n_L * n_rb);
    LBG_CUDA(cudaMemcpyAsync(c->n0_L.p, L, (size_t)n_L * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, l1, (size_t)n_l1 * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, w_l1, (size_t)n_l1 * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->n0_rowflag.p, rowflag.data(), (size_t)n_l1 * 4, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d_ph, ph.data(), ph.size() * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d_phf, phf.data(), phf.size() * 8, cudaMemcpyHostToDevice, st));
    if (n_special) {
        LBG_CUDA(cudaMemcpyAsync(c->s3.p, snu.data(), snu_bytes, cudaMemcpyHostToDevice, st));
        LBG_CUDA(cudaMemcpyAsync(c->s4.p, srow.data(), (size_t)n_special * 4, cudaMemcpyHostToDevice, st));
        if (n_exact)
            LBG_CUDA(cudaMemcpyAsync(c->s2.p, exact_u, (size_t)n_exact * n_phi * 64, cudaMemcpyHostToDevice, st));
    }

    if (variant >= 1) LBG_CUDA(cudaMemcpyToSymbolAsync(c_phi, ph.data(), ph.size() * 8, 0, cudaMemcpyHostToDevice, st));
    KernelTimer t(c);
    n0_uside_kernel<<<(n_l1 + 127) / 128, 128, 0, st>>>(P, c->s0.as<double>(), c->s1.as<double>(), c->n0_rowflag.as<int>(),
                                                         c->cl_rows.as<ClRow>(), c->n0_utab.as<URow>());
    t.launched();
    if (n_special > n_exact) {   // library-evaluated splines at |u|_j for the rows the caller gave no values for
        int cnt = (n_special - n_exact) * n_phi;
        n0_special_u_kernel<<<(cnt + 127) / 128, 128, 0, st>>>(c->n_ell, cnt, c->s3.as<double>() + (size_t)n_exact * n_phi,
                                                               c->cl_rows.as<ClRow>(), c->s2.as<double>() + (size_t)n_exact * n_phi * 8);
        t.launched();
    }
    if (variant == 3) {
        std::vector<double> cw(2 * (size_t)n_phe);
        for (int j = 0; j < n_phe; ++j) { cw[2 * j] = ph[j]; cw[2 * j + 1] = ph[2 * (size_t)n_phe + j]; }
        LBG_CUDA(cudaMemcpyToSymbolAsync(c_cw, cw.data(), cw.size() * 8, 0, cudaMemcpyHostToDevice, st));
        LBG_CUDA(cudaStreamSynchronize(st));      // cw is a local
        const int n_int = c->n_ell - 1;
        n0_poly_table_kernel<<<(n_int + 127) / 128, 128, 0, st>>>(P, fit, n_int, c->cl_soa.as<double>(), c->cl_poly.as<double4>());
        t.launched();
        cudaEventRecord(c->ev2, st);
        const bool full = terms_mask == (LBG_TT | LBG_TE | LBG_EE | LBG_EB);
        const dim3 grid(n_rb, n_L);
#define LBG_R3(UB, FULL) n0_rows3_kernel<UB, FULL><<<grid, 128, 0, st>>>(P, c->n0_L.as<double>(), c->n0_utab.as<URow>(), \
                             c->cl_poly.as<double4>(), c->cl_soa.as<double>(), d_part)
#define LBG_R5(UB, FULL) n0_rows5_kernel<UB, FULL><<<grid, 128, 0, st>>>(P, c->n0_L.as<double>(), c->n0_utab.as<URow>(), \
                             c->cl_poly.as<double4>(), c->cl_soa.as<double>(), d_part)
        if (rows5) {
            if (c->ub_zero) { if (full) LBG_R5(false, true); else LBG_R5(false, false); }
            else            { if (full) LBG_R5(true, true);  else LBG_R5(true, false); }
        } else {
            if (c->ub_zero) { if (full) LBG_R3(false, true); else LBG_R3(false, false); }
            else            { if (full) LBG_R3(true, true);  else LBG_R3(true, false); }
        }
#undef LBG_R5
#undef LBG_R3
    } else if (variant == 2) {
        // overlapping-pair copy of the spline coefficients, built from the spectrum-major table (only this variant reads it)
        if (c->cl_pairs.reserve((size_t)8 * LBG_PAIR_STRIDE * 16)) return LBG_ENOMEM;
        LBG_CUDA(cudaMemsetAsync(c->cl_pairs.p, 0, (size_t)8 * LBG_PAIR_STRIDE * 16, st));
        n0_pairs_kernel<<<dim3((c->n_ell + 2 + 255) / 256, 8), 256, 0, st>>>(c->n_ell + 2, c->cl_soa_stride, c->cl_soa.as<double>(), c->cl_pairs.as<double>());
        t.launched();
        const double xmax = P.a * 2.0 * (double)c->n_ell;                 // largest Taylor argument of the noise factor
        const size_t smem = ((size_t)4 * n_phe + 5 * 4) * 8;
        n0_noise_knots_kernel<<<(c->n_ell + 255) / 256, 256, 0, st>>>(c->n_ell, P.a, c->n0_G.as<double>());
        t.launched();
        if (xmax < 0.03)
            n0_rows2_kernel<false><<<dim3(n_rb, n_L), 128, smem, st>>>(P, c->n0_L.as<double>(), c->n0_utab.as<URow>(), c->cl_pairs.as<double>(), c->cl_soa.as<double>(), c->n0_G.as<double>(), d_part);
        else
            n0_rows2_kernel<true><<<dim3(n_rb, n_L), 128, smem, st>>>(P, c->n0_L.as<double>(), c->n0_utab.as<URow>(), c->cl_pairs.as<double>(), c->cl_soa.as<double>(), c->n0_G.as<double>(), d_part);
    } else if (variant == 1) {
        n0_rows_kernel<<<dim3(n_rb, n_L), 128, 0, st>>>(P, c->n0_L.as<double>(), c->n0_utab.as<URow>(), c->cl_soa.as<ClRow>(), d_part);
    } else {
        n0_kernel<<<dim3(bx, n_chunk), 128, 0, st>>>(P, c->n0_L.as<double>(), d_ph, c->n0_utab.as<URow>(),
                                                     c->cl_rows.as<ClRow>(), d_part);
    }
    t.launched();
    cudaEventRecord(c->ev3, st);
    if (n_special) {
        n0_special_kernel<<<(int)((npairs_full + 127) / 128), 128, 0, st>>>(P, n_special, n_phi, c->s4.as<int>(), c->n0_L.as<double>(), d_phf,
                                                                            c->n0_utab.as<URow>(), c->s2.as<double>(), c->s3.as<double>(),
                                                                            c->cl_rows.as<ClRow>(), d_spart);
        t.launched();
    }
    if (variant >= 1)
        n0_rows_reduce_kernel<<<n_L, 128, 0, st>>>(P, n_rb, c->n0_L.as<double>(), d_part, n_special ? d_spart : nullptr, n_phi, c->n0_out.as<double>());
    else
        n0_reduce_kernel<<<n_L, 128, 0, st>>>(P, c->n0_L.as<double>(), d_part, n_special ? d_spart : nullptr, n_phi, c->n0_out.as<double>());
    t.launched();
    std::vector<double> host((size_t)6 * n_L);
    // Wait for the kernels here, not inside the copy: a device-to-host copy into pageable memory issued while the stream is busy
    // blocks in the driver for the whole kernel time and (measured, scripts/overlap_timeline.py) stalls the pageable copies of
    // every other context of the process with it, which would serialise lane-0 calls behind this 25 ms launch.
    LBG_CUDA(cudaStreamSynchronize(st));
    LBG_CUDA(cudaMemcpyAsync(host.data(), c->n0_out.p, host.size() * 8, cudaMemcpyDeviceToHost, st));
    int rc = t.finish();
    if (rc) return rc;
    { float ms = 0.f; c->last_main_ms = (variant == 3 && cudaEventElapsedTime(&ms, c->ev2, c->ev3) == cudaSuccess) ? ms : c->last_ms; }
    memcpy(out_nkk, host.data(), (size_t)n_L * 8);
    if (out_terms) memcpy(out_terms, host.data() + n_L, (size_t)5 * n_L * 8);
    if (nodes_executed) *nodes_executed = npairs * (active_rows - n_special) + npairs_full * n_special;
    return LBG_OK;
}
