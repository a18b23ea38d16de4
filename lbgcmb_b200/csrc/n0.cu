// n0.cu -- Hu-Okamoto flat-sky N0 reconstruction noise, root flavour (replaces cmb/Nkk.py:54-172).
//
// Work decomposition (B200: 148 SMs, FP64 pipe + L1 data pipe bound, tables L2-resident):
//   thread <-> one |u| = l1 row of one L; it loops over the phi nodes.  CTA = 128 consecutive rows, grid = n_L x 79 row blocks.
//   Everything that depends on l1 only (spectra at u, totals, reciprocals, weight * l1) is computed once per call into a
//   128-byte "u row" (n0_uside_kernel) and lives in registers for the whole phi loop.  The v-side spectra are cubic splines
//   evaluated by Horner from a per-call table of per-interval monomial coefficients with the detector noise folded in
//   (n0_poly_table_kernel).  Partial sums go to a scratch buffer and are reduced in a fixed order (deterministic, no atomics).
//
// Two integrand kernels, chosen by the inputs and reported by lbg_n0_last_path():
//   n0_rows5_kernel (path 5)  phi grid folded onto mirror pairs: |v|^2 is monotone in the node index, so every cut of
//                             prep_lmaxmask (cmb/Nkk.py:13-26) is an index RANGE found once per row; masked nodes are never visited.
//   n0_rows3_kernel (path 3)  the generic kernel: per-node masks.  Taken for an unfolded phi grid, for tables that end inside the
//                             reachable |v| range (the spline's fill value 0 takes over there, prep_camb.py:88-89), for beams so wide
//                             that the folded-noise fit leaves its 2e-12 bound (then exp() per node), or on request (fold_phi = 2).
// Both leave the nodes within 4e-6 of a cut in |v|^2 to a second pass that evaluates them with the reference's own operation order
// (n0_node), and both hand the few rows whose u side needs per-node treatment to n0_special_kernel.
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
    int n_L, n_phi, n_l1, n_ell, n_rb;
    int n_int;                    // knot intervals of the per-call polynomial table (covers every reachable |v|)
    int soa_stride;               // the spectrum-major B-spline table is double[8][soa_stride]
    int lane;                     // which block of c_cw holds this call's (cos phi_j, w_j)
    unsigned terms;
    double nT, nP, a;             // N_TT(l) = nT exp(a l (l+1)), N_EE = N_BB = nP exp(.)  (cmb/DetectorNoise.py:22-26)
    double lmax_T, lmax_P, lmin_v;
    double ebfac;
    int fold;                     // phi nodes are mirror pairs: phi[] holds [6][n_phi] cos,sin,w,cos_m,sin_m,frac_j
    unsigned near_lo[4], near_span[4];   // hi-word windows of |v|^2 around lmin_v^2, lmax_T^2, lmax_P^2, (last knot)^2
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

// One integrand node (cmb/Nkk.py:66-162) for |u| = l at angle (c, s) = (cos phi, sin phi) and lensing multipole L, in the
// reference's operation order.  actT / actP: the |u| side of prep_lmaxmask (cmb/Nkk.py:17,20) for the 3000 / 5000 families;
// wl = weight * l.  Returns whether the node passed a mask (it then counts as an unmasked node).
template <bool SOA = false>
__device__ __forceinline__ bool n0_node(const N0Params& P, const ClRow* __restrict__ rows, double L, double l, double c, double s,
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
    if (!(okT || okP)) return false;

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
    return true;
}

// (cos phi_j, w_j) of the (folded) phi grid, one block per context lane so that the two lanes of a GPU (lbg_ctx_create_lane) may
// run N0 calls with different grids at the same time.  Read through the uniform datapath: j is warp-uniform.
#define LBG_MAX_LANES 4
#define LBG_MAX_PHI 512
__constant__ double2 c_cw[LBG_MAX_LANES][LBG_MAX_PHI];

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

// ---- per-call polynomial table ----------------------------------------------------------------------------------------------
// What bounds the integrand kernels is not one pipe but three at once (ncu: FP64, L1 data pipe, issue): every coefficient a lane
// loads costs register-fill bandwidth (128 B/clk/SM, no dedupe between lanes).  They therefore
//  * evaluate each spectrum from the monomial coefficients of its knot interval (Horner, 3 DFMA, four 8-byte loads per slot),
//    coefficient-major with a compile-time stride so that neighbouring lanes (neighbouring |v|) find a coefficient in neighbouring bytes;
//  * fold the detector noise into the C^tot slots when the table is built for the call: on [i, i+1] the factor
//    exp(a t (2i+1+t)) - 1 is replaced by its cubic interpolant at the four Chebyshev nodes (relative error bounded by
//    (beta^4 + 12 a beta^2 + 12 a^2) / 3072, beta = a (2 vmax + 1); beyond 2e-12 the generic kernel calls exp() per node instead);
//  * never read the unlensed-BB slot when that spectrum is identically zero (CAMB scalar runs: prep_camb.py:84-85).
// Intervals beyond the last knot of the table hold the spline's fill value 0 (+ the noise): prep_camb.py:88-89.
struct ChebFit { double m[4][4]; double t[4]; };     // monomial coefficients = m . samples at t[]

// slots per interval: 0 C^EE_tot, 1 C^BB_tot, 2 unlensed EE, 3 unlensed BB | 4 C^TT_tot, 5 lensed TE, 6 unlensed TT, 7 unlensed TE
__global__ void n0_poly_table_kernel(N0Params P, ChebFit F, int fold_noise, const double* __restrict__ soa, double4* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P.n_int) return;
    const size_t st = (size_t)P.soa_stride;
    double4 a[8];
    const bool in_table = i < P.n_ell - 1;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        double c0 = 0.0, c1 = 0.0, c2 = 0.0, c3 = 0.0;
        if (in_table) { c0 = soa[k * st + i]; c1 = soa[k * st + i + 1]; c2 = soa[k * st + i + 2]; c3 = soa[k * st + i + 3]; }
        a[k].x = (c0 + 4.0 * c1 + c2) * (1.0 / 6.0);
        a[k].y = (c2 - c0) * 0.5;
        a[k].z = (c0 - 2.0 * c1 + c2) * 0.5;
        a[k].w = ((c3 - c0) + 3.0 * (c1 - c2)) * (1.0 / 6.0);
    }
    // fold_noise: 0 = no noise in the table (the generic kernel calls exp() per node), 1 = noise folded into the three C^tot slots,
    // 2 = C^tot slots hold the lensed splines alone and slot 8 holds the noise SHAPE exp(a l (l + 1)) (multi-level kernel:
    // C^tot = S + n_level G, so every level of one beam shares these loads)
    double n[4] = {0.0, 0.0, 0.0, 0.0};
    if (fold_noise) {
        const double fi = (double)i;
        double g[4];
#pragma unroll
        for (int q = 0; q < 4; ++q) g[q] = expm1((P.a * F.t[q]) * fma(2.0, fi, 1.0 + F.t[q]));
        const double G = exp((fi * (fi + 1.0)) * P.a);
#pragma unroll
        for (int q = 0; q < 4; ++q) n[q] = G * (F.m[q][0] * g[0] + F.m[q][1] * g[1] + F.m[q][2] * g[2] + F.m[q][3] * g[3]);
        n[0] += G;
    }
    // layout double[slot][coefficient][LBG_PAIR_STRIDE]: four 8-byte loads per slot.  scripts/micro/ldg_pattern.cu (B200): a warp
    // LDG.E.64 whose lane pairs share an element costs 2.6 SM-clk on the L1-hit path, an LDG.E.128 of the same shape 6.9 -- two 64-bit
    // loads move the same bytes 24 % cheaper, and the 16-17 intervals a warp touches are 136 contiguous bytes of one array
    double* o = reinterpret_cast<double*>(out) + i;
#define LBG_PUT(slot, v) o[(4 * (slot)) * LBG_PAIR_STRIDE] = (v).x; o[(4 * (slot) + 1) * LBG_PAIR_STRIDE] = (v).y; \
                         o[(4 * (slot) + 2) * LBG_PAIR_STRIDE] = (v).z; o[(4 * (slot) + 3) * LBG_PAIR_STRIDE] = (v).w
    const double nPf = fold_noise == 2 ? 0.0 : P.nP, nTf = fold_noise == 2 ? 0.0 : P.nT;
    const double4 e2 = make_double4(fma(nPf, n[0], a[6].x), fma(nPf, n[1], a[6].y), fma(nPf, n[2], a[6].z), fma(nPf, n[3], a[6].w));
    const double4 b2 = make_double4(fma(nPf, n[0], a[7].x), fma(nPf, n[1], a[7].y), fma(nPf, n[2], a[7].z), fma(nPf, n[3], a[7].w));
    const double4 t2 = make_double4(fma(nTf, n[0], a[4].x), fma(nTf, n[1], a[4].y), fma(nTf, n[2], a[4].z), fma(nTf, n[3], a[4].w));
    if (fold_noise == 2) { const double4 g4 = make_double4(n[0], n[1], n[2], n[3]); LBG_PUT(8, g4); }
    LBG_PUT(0, e2); LBG_PUT(1, b2); LBG_PUT(2, a[2]); LBG_PUT(3, a[3]);
    LBG_PUT(4, t2); LBG_PUT(5, a[5]); LBG_PUT(6, a[0]); LBG_PUT(7, a[1]);
#undef LBG_PUT
}

template <int SLOT>
__device__ __forceinline__ double horner3(const double* __restrict__ p, double t) {    // p = table + interval index
    const double a0 = __ldg(p + (4 * SLOT) * LBG_PAIR_STRIDE), a1 = __ldg(p + (4 * SLOT + 1) * LBG_PAIR_STRIDE);
    const double a2 = __ldg(p + (4 * SLOT + 2) * LBG_PAIR_STRIDE), a3 = __ldg(p + (4 * SLOT + 3) * LBG_PAIR_STRIDE);
    return fma(fma(fma(a3, t, a2), t, a1), t, a0);
}

struct RowGeom { double L2, l2, rl2x2, lL, m2lL, l2pL2; };   // per-thread constants of one (L, |u|) row

// Row of this thread + the per-CTA epilogue shared by both integrand kernels.
struct RowCtx { int fl; double l, wl; UVals u; };
__device__ __forceinline__ RowCtx load_row(const N0Params& P, const URow* __restrict__ utab, int r) {
    RowCtx c;
    c.fl = 0; c.l = 1.0; c.wl = 0.0; c.u = UVals{};
    if (r < P.n_l1) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q0 = ld4(up), q1 = ld4(up + 1), q2 = ld4(up + 2), q3 = ld4(up + 3);
        c.fl = (int)q3.y;
        c.l = q0.x; c.wl = q0.y;
        c.u.uT = q0.z; c.u.uX = q0.w; c.u.uE = q1.x; c.u.uB = q1.y; c.u.T1 = q1.z; c.u.E1 = q1.w;
        c.u.X1 = q2.x; c.u.rT1h = q2.y; c.u.rE1h = q2.z; c.u.rE1 = q2.w; c.u.rB1h = q3.x;
    }
    return c;
}

// partial[6][n_L][n_rb]: the five estimator sums of this CTA's 128 rows and, in slot 5, the number of unmasked nodes it evaluated.
__device__ __forceinline__ void store_partials(const N0Params& P, double (*red)[4], int iL, int rb, double wl, double aTT, double aTE,
                                               double aEE, double aEB, double aBB, double cnt, double* __restrict__ partial) {
    double acc[6] = {aTT * wl, aTE * wl, aEE * wl, aEB * wl, aBB * wl, cnt};
#pragma unroll
    for (int t = 0; t < 6; ++t) {
        double v = acc[t];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0) red[t][threadIdx.x >> 5] = v;
    }
    __syncthreads();
    if (threadIdx.x < 6) {
        const int t = threadIdx.x;
        const double v = (red[t][0] + red[t][1]) + (red[t][2] + red[t][3]);
        partial[((size_t)t * P.n_L + iL) * P.n_rb + rb] = v;
    }
}

// ---- generic integrand kernel: per-node masks (path 3) --------------------------------------------------------------------------
// EXPN: the detector noise is NOT in the table; C^tot(v) = spline + N exp(a |v| (|v| + 1)) per node.
template <bool DO_T, bool EXPN>
__device__ __forceinline__ void n0_poly_node(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                             bool actT, bool actP, const double2 cw, bool& any_rare, double& cnt,
                                             double& aTT, double& aTE, double& aEE, double& aEB, double& aBB) {
    const double nv2 = fma(g.m2lL, cw.x, g.l2pL2);               // |v|^2 = l^2 + L^2 - 2 l L cos phi
    const unsigned hi = (unsigned)__double2hiint(nv2);
    const bool rare = (hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] ||
                      (hi - P.near_lo[2]) <= P.near_span[2] || (hi - P.near_lo[3]) <= P.near_span[3];
    any_rare = any_rare || rare;                                 // the bit-faithful second pass only runs for rows that saw one
    const bool vok = !rare && hi > P.near_lo[0];
    const bool okT = DO_T && actT && vok && hi < P.near_lo[1];
    const bool okP = actP && vok && hi < P.near_lo[2];
    if (!__any_sync(0xffffffffu, okT || okP)) return;
    const bool doT = DO_T && __any_sync(0xffffffffu, okT);       // warp-uniform
    cnt += (okT || okP) ? 1.0 : 0.0;

    const double y = rsqrt_fast(nv2);                            // lanes with a bad |v|^2 carry NaN; their terms are selected away
    const double nv = nv2 * y;
    const double dLu = g.lL * cw.x, dLv = g.L2 - dLu;            // L.u, L.v
    const double dot = dLu - g.l2;                               // u.v
    const double c2 = fma((dot * dot) * (y * y), g.rl2x2, -1.0); // cos 2phi_uv
    const double s2sq = fma(-c2, c2, 1.0);                       // sin^2 2phi_uv

    const int i = max(min((int)nv, P.n_int - 1), 0);
    const double t = nv - (double)i;
    const double* row = reinterpret_cast<const double*>(ptab) + i;
    double E2 = horner3<0>(row, t);
    const double vE = horner3<2>(row, t);
    double B2 = horner3<1>(row, t);
    const double vB = horner3<3>(row, t);
    double en = 0.0;
    if (EXPN) {
        en = exp((nv * (nv + 1.0)) * P.a);
        E2 = fma(P.nP, en, E2);
        B2 = fma(P.nP, en, B2);
    }
    if (!(P.terms & (LBG_EB | LBG_BB))) B2 = 1.0;
    const double uEd = u.uE * dLu;
    const double p2 = E2 * B2;
    double rE2, rB2;
    if (doT) {
        double T2 = horner3<4>(row, t);
        if (EXPN) T2 = fma(P.nT, en, T2);
        const double X2 = horner3<5>(row, t), vT = horner3<6>(row, t), vX = horner3<7>(row, t);
        const double E1T2 = u.E1 * T2, xx = u.X1 * X2;
        double den = fma(u.T1 * E2, E1T2, -(xx * xx));
        if (!(P.terms & LBG_TE)) den = 1.0;
        if (!(P.terms & LBG_TT)) T2 = 1.0;
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
    const double fb = fma(-vB, dLv, uEd);
    const double eb = ((fb * fb) * s2sq) * (u.rE1 * rB2);
    aEE = fma(cw.y, okP ? ee : 0.0, aEE);
    aEB = fma(cw.y, okP ? eb : 0.0, aEB);
    if (P.terms & LBG_BB) {                                      // cmb/Nkk.py:148-154
        const double fbb = fma(vB, dLv, u.uB * dLu) * c2;
        const double bb = (fbb * fbb) * (u.rB1h * rB2);
        aBB = fma(cw.y, okP ? bb : 0.0, aBB);
    }
}

template <bool DO_T, bool EXPN>
__device__ __forceinline__ void n0_rows3_loop(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                              bool actT, bool actP, bool& any_rare, double& cnt,
                                              double& aTT, double& aTE, double& aEE, double& aEB, double& aBB) {
    const double2* cw = c_cw[P.lane];
    int j = 0;
    for (; j + 1 < P.n_phi; j += 2) {      // two independent nodes per trip: ILP for the FP64 pipe
        n0_poly_node<DO_T, EXPN>(P, ptab, g, u, actT, actP, cw[j], any_rare, cnt, aTT, aTE, aEE, aEB, aBB);
        n0_poly_node<DO_T, EXPN>(P, ptab, g, u, actT, actP, cw[j + 1], any_rare, cnt, aTT, aTE, aEE, aEB, aBB);
    }
    for (; j < P.n_phi; ++j) n0_poly_node<DO_T, EXPN>(P, ptab, g, u, actT, actP, cw[j], any_rare, cnt, aTT, aTE, aEE, aEB, aBB);
}

template <bool EXPN>
__global__ void __launch_bounds__(128, 4)
n0_rows3_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const double4* __restrict__ ptab,
                const double* __restrict__ soa, const double* __restrict__ phi /*[6][n_phi]*/, double* __restrict__ partial) {
    __shared__ double red[6][4];
    __shared__ int s_anyT;
    const int iL = blockIdx.x / P.n_rb, rb = blockIdx.x - iL * P.n_rb;
    const int r = rb * 128 + threadIdx.x;
    const double L = Ls[iL];
    if (threadIdx.x == 0) s_anyT = 0;
    const RowCtx rc = load_row(P, utab, r);
    const int fl = rc.fl;
    const double l = rc.l;
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    __syncthreads();
    if (actT) s_anyT = 1;                                           // benign race: every writer stores 1
    __syncthreads();
    RowGeom g;
    g.L2 = L * L; g.l2 = l * l; g.rl2x2 = 2.0 / g.l2; g.lL = l * L; g.m2lL = -2.0 * g.lL; g.l2pL2 = g.l2 + g.L2;
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0, cnt = 0.0;
    bool any_rare = false;
    if (s_anyT) n0_rows3_loop<true, EXPN>(P, ptab, g, rc.u, actT, actP, any_rare, cnt, aTT, aTE, aEE, aEB, aBB);
    else        n0_rows3_loop<false, EXPN>(P, ptab, g, rc.u, actT, actP, any_rare, cnt, aTT, aTE, aEE, aEB, aBB);
    // second pass: nodes within 4e-6 of a cut in |v|^2, evaluated exactly as the reference rounds them
    if (act && any_rare) {
        const bool fold = P.fold != 0;
        const int n = P.n_phi;
        const double2* cw = c_cw[P.lane];
        for (int j = 0; j < n; ++j) {
            const double nv2 = fma(g.m2lL, cw[j].x, g.l2pL2);
            const unsigned hi = (unsigned)__double2hiint(nv2);
            if ((hi - P.near_lo[0]) <= P.near_span[0] || (hi - P.near_lo[1]) <= P.near_span[1] ||
                (hi - P.near_lo[2]) <= P.near_span[2] || (hi - P.near_lo[3]) <= P.near_span[3])
                cnt += n0_node<true>(P, reinterpret_cast<const ClRow*>(soa), L, l, __ldg(phi + j), __ldg(phi + n + j), rc.u, fl & 1, fl & 4,
                                     cw[j].y, fold, __ldg(phi + 3 * n + j), __ldg(phi + 4 * n + j), __ldg(phi + 5 * n + j),
                                     aTT, aTE, aEE, aEB, aBB) ? 1.0 : 0.0;
        }
    }
    if (!(P.terms & LBG_TT)) aTT = 0.0;
    if (!(P.terms & LBG_TE)) aTE = 0.0;
    if (!(P.terms & LBG_EE)) aEE = 0.0;
    if (!(P.terms & LBG_EB)) aEB = 0.0;
    store_partials(P, red, iL, rb, rc.wl, aTT, aTE, aEE, aEB, aBB, cnt, partial);
}

// ---- fast integrand kernel: masks turned into phi-index ranges (path 5) ---------------------------------------------------------
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
__device__ __forceinline__ int first_j_hi_ge(const RowGeom& g, const double2* cw, int n_phi, unsigned x) {   // first j with hi(j) >= x (n_phi if none)
    int lo = 0, hi = n_phi;
    while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const unsigned h = (unsigned)__double2hiint(fma(g.m2lL, cw[mid].x, g.l2pL2));
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
    q.i = max(min((int)nv, P.n_int - 1), 0);
    q.t = nv - (double)q.i;
    return q;
}

template <bool DO_T, bool HAS_UB, bool FULL>
__device__ __forceinline__ void n0_node5(const N0Params& P, const double4* __restrict__ ptab, const RowGeom& g, const UVals& u,
                                         bool okT, bool okP, const double2 cw, const NodePre pre,
                                         double& aTT, double& aTE, double& aEE, double& aEB, double& aBB) {
    const double y = pre.y, t = pre.t;
    const double* row = reinterpret_cast<const double*>(ptab) + pre.i;
    const double dLu = g.lL * cw.x, dLv = g.L2 - dLu;            // L.u, L.v
    const double dot = dLu - g.l2;                               // u.v
    const double c2 = fma((dot * dot) * (y * y), g.rl2x2, -1.0); // cos 2phi_uv = 2 (u.v)^2 / (|u|^2 |v|^2) - 1
    const double s2sq = fma(-c2, c2, 1.0);                       // sin^2 2phi_uv
    const double E2 = horner3<0>(row, t), vE = horner3<2>(row, t);
    double B2 = horner3<1>(row, t);
    const double vB = HAS_UB ? horner3<3>(row, t) : 0.0;
    if (!FULL) { if (!(P.terms & (LBG_EB | LBG_BB))) B2 = 1.0; }
    const double uEd = u.uE * dLu;
    const double p2 = E2 * B2;
    double rE2, rB2;
    if (DO_T) {
        double T2 = horner3<4>(row, t);
        const double X2 = horner3<5>(row, t), vT = horner3<6>(row, t), vX = horner3<7>(row, t);
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
    if (!FULL && HAS_UB) {
        if (P.terms & LBG_BB) {                                  // cmb/Nkk.py:148-154 (zero when the unlensed BB table is)
            const double fbb = fma(vB, dLv, u.uB * dLu) * c2;
            const double bb = (fbb * fbb) * (u.rB1h * rB2);
            aBB = fma(cw.y, okP ? bb : 0.0, aBB);
        }
    }
}

template <bool HAS_UB, bool FULL>
__global__ void __launch_bounds__(128, 4)
n0_rows5_kernel(N0Params P, const double* __restrict__ Ls, const URow* __restrict__ utab, const double4* __restrict__ ptab,
                const double* __restrict__ soa, const double* __restrict__ phi /*[6][n_phi]*/, double* __restrict__ partial) {
    __shared__ double red[6][4];
    const int iL = blockIdx.x / P.n_rb, rb = blockIdx.x - iL * P.n_rb;
    const int r = rb * 128 + threadIdx.x;
    const double L = Ls[iL];
    const RowCtx rc = load_row(P, utab, r);
    const int fl = rc.fl;
    const double l = rc.l;
    const UVals& u = rc.u;
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    RowGeom g;
    g.L2 = L * L; g.l2 = l * l; g.rl2x2 = 2.0 / g.l2; g.lL = l * L; g.m2lL = -2.0 * g.lL; g.l2pL2 = g.l2 + g.L2;
    const int n = P.n_phi;
    const double2* cw = c_cw[P.lane];
    // range ends of this row (see the header); inactive lanes get empty ranges that do not widen the warp's walk
    int j0s = n, j0e = n, jT = 0, jTe = 0, jP = 0, jPe = 0;
    if (act) {
        j0s = first_j_hi_ge(g, cw, n, P.near_lo[0]);
        j0e = first_j_hi_ge(g, cw, n, P.near_lo[0] + P.near_span[0] + 1u);
        jT = first_j_hi_ge(g, cw, n, P.near_lo[1]);
        jTe = first_j_hi_ge(g, cw, n, P.near_lo[1] + P.near_span[1] + 1u);
        jP = first_j_hi_ge(g, cw, n, P.near_lo[2]);
        jPe = first_j_hi_ge(g, cw, n, P.near_lo[2] + P.near_span[2] + 1u);
    }
    const bool any_rare = act && (j0e > j0s || jTe > jT || jPe > jP);
    const int endT = actT ? jT : 0, endP = actP ? jP : 0;       // first j at which the T / P estimators stop (|v| side)
    const int w0 = __reduce_min_sync(0xffffffffu, j0e);
    const int wT = __reduce_max_sync(0xffffffffu, endT), wP = __reduce_max_sync(0xffffffffu, endP);
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;
    // unmasked nodes of this row on the fast path: the P range [j0e, endP) without the lmax_T window (T nodes are P nodes too: lmax_T <= lmax_P)
    int cnt_i = max(endP - j0e, 0) - max(min(jTe, endP) - max(jT, j0e), 0);
    if (cnt_i < 0) cnt_i = 0;
#define LBG_OKT(j) ((j) >= j0e && (j) < endT)
#define LBG_OKP(j) ((j) >= j0e && (j) < endP && !((j) >= jT && (j) < jTe))
    // pairs (j, j+1) from w0 on: 7-slot body while j < wT, then 3-slot body while j < wP.  A pair's second node may lie past a
    // range end (or past n_phi): its predicates are false there, and c_cw holds finite entries up to n_phi + 3.
    int j = w0;
    NodePre p0 = n0_prep5(P, g, cw[j].x), p1 = n0_prep5(P, g, cw[j + 1].x);
    for (; j < wT; j += 2) {
        const NodePre n0 = n0_prep5(P, g, cw[j + 2].x), n1 = n0_prep5(P, g, cw[j + 3].x);
        n0_node5<true, HAS_UB, FULL>(P, ptab, g, u, LBG_OKT(j), LBG_OKP(j), cw[j], p0, aTT, aTE, aEE, aEB, aBB);
        n0_node5<true, HAS_UB, FULL>(P, ptab, g, u, LBG_OKT(j + 1), LBG_OKP(j + 1), cw[j + 1], p1, aTT, aTE, aEE, aEB, aBB);
        p0 = n0; p1 = n1;
    }
    for (; j < wP; j += 2) {
        const NodePre n0 = n0_prep5(P, g, cw[j + 2].x), n1 = n0_prep5(P, g, cw[j + 3].x);
        n0_node5<false, HAS_UB, FULL>(P, ptab, g, u, false, LBG_OKP(j), cw[j], p0, aTT, aTE, aEE, aEB, aBB);
        n0_node5<false, HAS_UB, FULL>(P, ptab, g, u, false, LBG_OKP(j + 1), cw[j + 1], p1, aTT, aTE, aEE, aEB, aBB);
        p0 = n0; p1 = n1;
    }
#undef LBG_OKT
#undef LBG_OKP
    double cnt = (double)cnt_i;
    if (any_rare) {                                              // bit-faithful second pass over the window nodes only
        const bool fold = P.fold != 0;
        for (int k = 0; k < 3; ++k) {
            const int a = k == 0 ? j0s : (k == 1 ? jT : jP), b = k == 0 ? j0e : (k == 1 ? jTe : jPe);
            for (int jj = a; jj < b; ++jj) {
                if (k == 2 && jj >= jT && jj < jTe) continue;    // lmax_T == lmax_P: the windows coincide, visit once
                if (k == 1 && jj >= j0s && jj < j0e) continue;
                if (k == 2 && jj >= j0s && jj < j0e) continue;
                cnt += n0_node<true>(P, reinterpret_cast<const ClRow*>(soa), L, l, __ldg(phi + jj), __ldg(phi + n + jj), u, fl & 1, fl & 4,
                                     cw[jj].y, fold, __ldg(phi + 3 * n + jj), __ldg(phi + 4 * n + jj), __ldg(phi + 5 * n + jj),
                                     aTT, aTE, aEE, aEB, aBB) ? 1.0 : 0.0;
            }
        }
    }
    if (!FULL) {
        if (!(P.terms & LBG_TT)) aTT = 0.0;
        if (!(P.terms & LBG_TE)) aTE = 0.0;
        if (!(P.terms & LBG_EE)) aEE = 0.0;
        if (!(P.terms & LBG_EB)) aEB = 0.0;
    }
    store_partials(P, red, iL, rb, rc.wl, aTT, aTE, aEE, aEB, aBB, cnt, partial);
}


// ---- several noise levels of ONE beam per node (survey-design sweeps: BASELINE configs[4], snr.py:108-124 loops Nkk over experiments) ------
// For levels that share the beam, C^tot_X(l) = S_X(l) + n_X^level G(l) with S the lensed spline and G(l) = exp(a l (l + 1)) the noise
// shape: the table holds S_TT, S_EE, S_BB and G once (n0_poly_table_kernel, fold_noise = 2) and a node serves NLEV levels with 8 slot
// reads (32 loads) + 3 FMAs per level instead of 7 slot reads per level -- the single-level kernel is bound by exactly those loads (L1
// data pipe 79 % against FP64 53 %, profiles/r2a_n0_rows5_ncu_full_summary.txt).  Geometry, masks (they do not depend on the noise), the
// four unlensed spectra and every level-independent product (L.u, L.v, cos 2phi, the f_alpha) are computed once per node.
// A level's arithmetic does not depend on NLEV or on its position in the group: its result is the same in any grouping.  It differs
// from n0_rows5_kernel's in the last bits only (noise added after the Horner sums instead of inside the coefficients).
struct LevelNoise { double nT, nP; };
#define LBG_MAX_NLEV 3
struct LevelParams { LevelNoise lv[LBG_MAX_NLEV]; };

template <int NLEV>
struct LevU { double T1[NLEV], E1[NLEV], rT1h[NLEV], rE1h[NLEV], rE1[NLEV]; };

template <int NLEV, bool DO_T, bool HAS_UB>
__device__ __forceinline__ void n0_node5m(const double4* __restrict__ ptab, const RowGeom& g, const UVals& u, const LevU<NLEV>& lu,
                                          const LevelParams& LP, bool okT, bool okP, const double2 cw, const NodePre pre,
                                          double (&aTT)[NLEV], double (&aTE)[NLEV], double (&aEE)[NLEV], double (&aEB)[NLEV]) {
    const double y = pre.y, t = pre.t;
    const double* row = reinterpret_cast<const double*>(ptab) + pre.i;
    const double dLu = g.lL * cw.x, dLv = g.L2 - dLu;
    const double dot = dLu - g.l2;
    const double c2 = fma((dot * dot) * (y * y), g.rl2x2, -1.0);
    const double s2sq = fma(-c2, c2, 1.0);
    const double SE = horner3<0>(row, t), vE = horner3<2>(row, t), SB = horner3<1>(row, t), G = horner3<8>(row, t);
    const double vB = HAS_UB ? horner3<3>(row, t) : 0.0;
    const double uEd = u.uE * dLu;
    const double fe = fma(vE, dLv, uEd) * c2;
    const double fe2 = fe * fe;
    const double fb = HAS_UB ? fma(-vB, dLv, uEd) : uEd;
    const double fb2s = (fb * fb) * s2sq;
    double ST = 0.0, X2 = 0.0, fa2 = 0.0, fa_uv = 0.0, fa_vu = 0.0, xx = 0.0;
    if (DO_T) {
        ST = horner3<4>(row, t); X2 = horner3<5>(row, t);
        const double vT = horner3<6>(row, t), vX = horner3<7>(row, t);
        xx = u.X1 * X2;
        const double fa = fma(vT, dLv, u.uT * dLu);
        fa2 = fa * fa;
        const double uXd = u.uX * dLu, vXd = vX * dLv;
        fa_uv = fma(uXd, c2, vXd); fa_vu = fma(vXd, c2, uXd);
    }
    const double wT = okT ? cw.y : 0.0, wP = okP ? cw.y : 0.0;     // a masked node adds +0 (its terms may be NaN: select BEFORE the product)
#pragma unroll
    for (int v = 0; v < NLEV; ++v) {
        const double E2 = fma(LP.lv[v].nP, G, SE), B2 = fma(LP.lv[v].nP, G, SB);
        const double p2 = E2 * B2;
        double rE2, rB2;
        if (DO_T) {
            const double T2 = fma(LP.lv[v].nT, G, ST);
            const double E1T2 = lu.E1[v] * T2;
            const double den = fma(lu.T1[v] * E2, E1T2, -(xx * xx));
            const double p1 = T2 * den;
            const double r = rcp_fast(p1 * p2);
            const double r34 = r * p1, r12 = r * p2;
            rE2 = r34 * B2; rB2 = r34 * E2;
            const double rT2 = r12 * den, rdn = r12 * T2;
            const double tt = fa2 * (lu.rT1h[v] * rT2);
            const double num = fma(E1T2, fa_uv, -(xx * fa_vu));
            const double te = (fa_uv * num) * rdn;
            aTT[v] = fma(wT, okT ? tt : 0.0, aTT[v]);
            aTE[v] = fma(wT, okT ? te : 0.0, aTE[v]);
        } else {
            const double r = rcp_fast(p2);
            rE2 = r * B2; rB2 = r * E2;
        }
        const double ee = fe2 * (lu.rE1h[v] * rE2);
        const double eb = fb2s * (lu.rE1[v] * rB2);
        aEE[v] = fma(wP, okP ? ee : 0.0, aEE[v]);
        aEB[v] = fma(wP, okP ? eb : 0.0, aEB[v]);
    }
}

template <int NLEV, bool HAS_UB>
__global__ void __launch_bounds__(128, 3)
n0_rows5m_kernel(N0Params P, LevelParams LP, const double* __restrict__ Ls, const URow* __restrict__ utab /*[NLEV][n_l1]*/,
                 const double4* __restrict__ ptab, const double* __restrict__ soa, const double* __restrict__ phi /*[6][n_phi]*/,
                 double* __restrict__ partial /*[NLEV][6][n_L][n_rb]*/) {
    __shared__ double red[6][4];
    const int iL = blockIdx.x / P.n_rb, rb = blockIdx.x - iL * P.n_rb;
    const int r = rb * 128 + threadIdx.x;
    const double L = Ls[iL];
    const RowCtx rc = load_row(P, utab, r);                      // level 0's row: everything level-independent comes from it
    const int fl = rc.fl;
    const double l = rc.l;
    const UVals& u = rc.u;
    LevU<NLEV> lu;
#pragma unroll
    for (int v = 0; v < NLEV; ++v) {
        lu.T1[v] = lu.E1[v] = 1.0; lu.rT1h[v] = lu.rE1h[v] = lu.rE1[v] = 0.0;
        if (r < P.n_l1) {
            const URow* ur = utab + (size_t)v * P.n_l1 + r;
            lu.T1[v] = ur->T1; lu.E1[v] = ur->E1; lu.rT1h[v] = ur->rT1h; lu.rE1h[v] = ur->rE1h; lu.rE1[v] = ur->rE1;
        }
    }
    const bool act = (fl & 5) && !(fl & 16);
    const bool actT = act && (fl & 1), actP = act && (fl & 4);
    RowGeom g;
    g.L2 = L * L; g.l2 = l * l; g.rl2x2 = 2.0 / g.l2; g.lL = l * L; g.m2lL = -2.0 * g.lL; g.l2pL2 = g.l2 + g.L2;
    const int n = P.n_phi;
    const double2* cw = c_cw[P.lane];
    int j0s = n, j0e = n, jT = 0, jTe = 0, jP = 0, jPe = 0;
    if (act) {
        j0s = first_j_hi_ge(g, cw, n, P.near_lo[0]);
        j0e = first_j_hi_ge(g, cw, n, P.near_lo[0] + P.near_span[0] + 1u);
        jT = first_j_hi_ge(g, cw, n, P.near_lo[1]);
        jTe = first_j_hi_ge(g, cw, n, P.near_lo[1] + P.near_span[1] + 1u);
        jP = first_j_hi_ge(g, cw, n, P.near_lo[2]);
        jPe = first_j_hi_ge(g, cw, n, P.near_lo[2] + P.near_span[2] + 1u);
    }
    const bool any_rare = act && (j0e > j0s || jTe > jT || jPe > jP);
    const int endT = actT ? jT : 0, endP = actP ? jP : 0;
    const int w0 = __reduce_min_sync(0xffffffffu, j0e);
    const int wT = __reduce_max_sync(0xffffffffu, endT), wP = __reduce_max_sync(0xffffffffu, endP);
    double aTT[NLEV], aTE[NLEV], aEE[NLEV], aEB[NLEV], aBB[NLEV];
#pragma unroll
    for (int v = 0; v < NLEV; ++v) aTT[v] = aTE[v] = aEE[v] = aEB[v] = aBB[v] = 0.0;
    int cnt_i = max(endP - j0e, 0) - max(min(jTe, endP) - max(jT, j0e), 0);
    if (cnt_i < 0) cnt_i = 0;
#define LBG_OKT(j) ((j) >= j0e && (j) < endT)
#define LBG_OKP(j) ((j) >= j0e && (j) < endP && !((j) >= jT && (j) < jTe))
    int j = w0;
    NodePre p0 = n0_prep5(P, g, cw[j].x);
    for (; j < wT; ++j) {                                        // one node per trip: the NLEV levels are the independent work
        const NodePre nx = n0_prep5(P, g, cw[j + 1].x);
        n0_node5m<NLEV, true, HAS_UB>(ptab, g, u, lu, LP, LBG_OKT(j), LBG_OKP(j), cw[j], p0, aTT, aTE, aEE, aEB);
        p0 = nx;
    }
    for (; j < wP; ++j) {
        const NodePre nx = n0_prep5(P, g, cw[j + 1].x);
        n0_node5m<NLEV, false, HAS_UB>(ptab, g, u, lu, LP, false, LBG_OKP(j), cw[j], p0, aTT, aTE, aEE, aEB);
        p0 = nx;
    }
#undef LBG_OKT
#undef LBG_OKP
    double cnt = (double)cnt_i;
    if (any_rare) {                                              // bit-faithful second pass over the window nodes, level by level
        const bool fold = P.fold != 0;
        for (int k = 0; k < 3; ++k) {
            const int a = k == 0 ? j0s : (k == 1 ? jT : jP), b = k == 0 ? j0e : (k == 1 ? jTe : jPe);
            for (int jj = a; jj < b; ++jj) {
                if (k == 2 && jj >= jT && jj < jTe) continue;
                if (k == 1 && jj >= j0s && jj < j0e) continue;
                if (k == 2 && jj >= j0s && jj < j0e) continue;
#pragma unroll
                for (int v = 0; v < NLEV; ++v) {
                    N0Params Pv = P;
                    Pv.nT = LP.lv[v].nT; Pv.nP = LP.lv[v].nP;
                    UVals uv = u;
                    uv.T1 = lu.T1[v]; uv.E1 = lu.E1[v]; uv.rT1h = lu.rT1h[v]; uv.rE1h = lu.rE1h[v]; uv.rE1 = lu.rE1[v];
                    const bool hit = n0_node<true>(Pv, reinterpret_cast<const ClRow*>(soa), L, l, __ldg(phi + jj), __ldg(phi + n + jj), uv, fl & 1, fl & 4,
                                                   cw[jj].y, fold, __ldg(phi + 3 * n + jj), __ldg(phi + 4 * n + jj), __ldg(phi + 5 * n + jj),
                                                   aTT[v], aTE[v], aEE[v], aEB[v], aBB[v]);
                    if (v == 0) cnt += hit ? 1.0 : 0.0;
                }
            }
        }
    }
    const size_t lev_stride = (size_t)6 * P.n_L * P.n_rb;
#pragma unroll
    for (int v = 0; v < NLEV; ++v) {
        if (v) __syncthreads();                                   // `red` is reused
        store_partials(P, red, iL, rb, rc.wl, aTT[v], aTE[v], aEE[v], aEB[v], aBB[v], cnt, partial + v * lev_stride);
    }
}

// Special rows, always on the full (unfolded) phi grid:
//  * rows sitting on the lmax_T / lmax_P cuts, where the reference's mask `|u| > lmax` (cmb/Nkk.py:17,20) depends on the
//    rounding of |u| = sqrt((l cos)^2 + (l sin)^2) at each phi node; rows on the last knot of a table that ends in range likewise;
//  * rows below the first non-zero C_l, where the reference's filter 1/C^tot is pure noise and its result inherits the rounding of
//    scipy's spline: their u-side values come from splev_exact() (a host restatement of that routine) or from the caller.
// su: [n_special][n_phi][8] splines at |u|_j (uTT,uTE,uEE,uBB,lTT,lTE,lEE,lBB); snu: [n_special][n_phi] |u|_j.
__global__ void __launch_bounds__(128)
n0_special_kernel(N0Params P, int n_special, int n_phi_full, const int* __restrict__ srow, const double* __restrict__ Ls,
                  const double* __restrict__ phi /*[3][n_phi_full]*/, const URow* __restrict__ utab,
                  const double* __restrict__ su, const double* __restrict__ snu, const ClRow* __restrict__ rows,
                  double* __restrict__ partial /*[6][n_L*n_phi_full]*/) {
    const int npairs = P.n_L * n_phi_full;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= npairs) return;
    const int iL = p / n_phi_full, j = p - iL * n_phi_full;
    const double L = Ls[iL], c = phi[j], s = phi[n_phi_full + j];
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0, cnt = 0.0;
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
        cnt += n0_node(P, rows, L, ur.l, c, s, u, !(nu > P.lmax_T), !(nu > P.lmax_P), ur.wl, false, 0.0, 0.0, 1.0, aTT, aTE, aEE, aEB, aBB) ? 1.0 : 0.0;
    }
    const double wphi = phi[2 * n_phi_full + j];
    const size_t stride = (size_t)npairs;
    partial[p] = aTT * wphi;
    partial[stride + p] = aTE * wphi;
    partial[2 * stride + p] = aEE * wphi;
    partial[3 * stride + p] = aEB * wphi;
    partial[4 * stride + p] = aBB * wphi;
    partial[5 * stride + p] = cnt;
}

// Library-side u values for the special rows on a cut: the eight splines at |u|_j.
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

// One block per L: fixed-order sum of the row-block partials (+ the special rows), then cmb/Nkk.py:169-172.
// out[7][n_L]: N0, the TT, TE, EE, EB, BB integrals, and the number of unmasked integrand nodes of this L.
__global__ void __launch_bounds__(128)
n0_rows_reduce_kernel(N0Params P, const double* __restrict__ Ls, const double* __restrict__ partial,
                      const double* __restrict__ spartial, int n_phi_full, double* __restrict__ out) {
    __shared__ double sm[6][128];
    const int iL = blockIdx.x;
    double acc[6] = {0, 0, 0, 0, 0, 0};
    for (int b = threadIdx.x; b < P.n_rb; b += 128) {
#pragma unroll
        for (int t = 0; t < 6; ++t) acc[t] += partial[((size_t)t * P.n_L + iL) * P.n_rb + b];
    }
    if (spartial) {
        const size_t sstride = (size_t)P.n_L * n_phi_full;
        for (int j = threadIdx.x; j < n_phi_full; j += 128) {
#pragma unroll
            for (int t = 0; t < 6; ++t) acc[t] += spartial[t * sstride + (size_t)iL * n_phi_full + j];
        }
    }
#pragma unroll
    for (int t = 0; t < 6; ++t) sm[t][threadIdx.x] = acc[t];
    __syncthreads();
    for (int h = 64; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
#pragma unroll
            for (int t = 0; t < 6; ++t) sm[t][threadIdx.x] += sm[t][threadIdx.x + h];
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
        for (int t = 0; t < 6; ++t) out[(size_t)(t + 1) * P.n_L + iL] = sm[t][0];
    }
}

// scipy's interp1d(kind='cubic') at one point, in its own operation order (ctx.cu).  On rows below the first non-zero C_l the
// reference's 1/C^tot is pure detector noise and inherits the last-bit rounding of that routine (DESIGN.md "parity notes"): those few
// values (2 rows x 201 nodes x 8 spectra) are computed on the host with the same recurrences from the same LAPACK-order coefficients.
double scipy_bspline_eval(const double* t, const double* c, int n, double x);
}  // namespace lbg

using namespace lbg;

namespace {
// Everything of an N0 call that depends on the quadrature grids and cuts only: built once, kept on the context (host copies for the
// comparison, device copies for the kernels) and reused while the caller passes the same grids (a forecast loop always does).
struct N0Key {
    int n_l1 = 0, n_phi = 0, fold = 0, n_exact = 0, n_ell = 0;
    double lmax_T = 0, lmax_P = 0, lmin_v = 0, exact_below = -1;
    std::vector<double> l1, w_l1, cphi, sphi, wphi;
    std::vector<int> exact_rows;
    bool same(int nl1, const double* a, const double* b, int nphi, const double* c, const double* s, const double* w, double lt, double lp,
              double lv, int f, int ne, const int* er, int nell) const {
        return nl1 == n_l1 && nphi == n_phi && f == fold && ne == n_exact && nell == n_ell && lt == lmax_T && lp == lmax_P && lv == lmin_v &&
               !memcmp(a, l1.data(), (size_t)nl1 * 8) && !memcmp(b, w_l1.data(), (size_t)nl1 * 8) && !memcmp(c, cphi.data(), (size_t)nphi * 8) &&
               !memcmp(s, sphi.data(), (size_t)nphi * 8) && !memcmp(w, wphi.data(), (size_t)nphi * 8) &&
               (ne == 0 || !memcmp(er, exact_rows.data(), (size_t)ne * 4));
    }
};
}  // namespace

struct lbg_n0_plan {
    N0Key key;
    bool valid = false;
    int n_phe = 0, n_special = 0, n_lib_exact = 0;   // folded node count; special rows; how many of them get library splev values
    long long active_rows = 0;
    bool cos_decreasing = false;
    std::vector<int> srow;
    std::vector<double> snu;                          // [n_special][n_phi] |u|_j in the reference's rounding
    std::vector<double> cw;                           // (cos, w) pairs of the main kernel's grid, zero-padded by 4
    double l1_max = 0.0;
};

// which context's (cos phi, w) block sits in lane `l` of c_cw on device `d` (c_cw is per device, contexts come and go)
static std::mutex g_cw_mu;
static lbg_ctx* g_cw_owner[64][LBG_MAX_LANES];

static void n0_plan_free(lbg_ctx* c) {
    delete c->n0_plan;
    c->n0_plan = nullptr;
    std::lock_guard<std::mutex> g(g_cw_mu);
    for (auto& dev : g_cw_owner) for (auto& o : dev) if (o == c) o = nullptr;
}

// ---- begin: validate, (re)build the plan, stage the inputs, enqueue every kernel and the result copy; returns without waiting ----
static int n0_begin_locked(lbg_ctx* c, int n_L, const double* L, unsigned terms_mask,
                           double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                           int n_l1, const double* l1, const double* w_l1,
                           int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                           double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                           int n_exact, const int* exact_rows, const double* exact_u, double exact_below,
                           int n_lev = 0, const double* deltaT_lev = nullptr, const int* iterative_lev = nullptr) {
    // n_lev > 0: several noise levels of this beam in one pass (lbg_n0_ho02_levels); deltaT_uK_arcmin / iterative are then ignored
    LBG_REQUIRE(c && L && l1 && w_l1 && cos_phi && sin_phi && w_phi, "lbg_n0_ho02: NULL argument");
    if (c->n_ell <= 0) { set_error("lbg_n0_ho02: call lbg_upload_cl_tables first"); return LBG_ESTATE; }
    if (c->n0_pending) { set_error("lbg_n0_ho02: a call begun with lbg_n0_ho02_begin is still pending on this context"); return LBG_ESTATE; }
    LBG_REQUIRE(n_L >= 0 && n_l1 > 0 && n_phi > 0, "lbg_n0_ho02: bad sizes n_L=%d n_l1=%d n_phi=%d", n_L, n_l1, n_phi);
    LBG_REQUIRE((terms_mask & ~(LBG_TT | LBG_TE | LBG_EE | LBG_EB | LBG_BB)) == 0 && terms_mask != 0,
                "lbg_n0_ho02: unsupported terms mask 0x%x", terms_mask);
    LBG_REQUIRE(Tcmb_K > 0 && lmax_T <= lmax_P, "lbg_n0_ho02: need Tcmb > 0 and lmax_T <= lmax_P");
    LBG_REQUIRE(lmin_v > 0.0 && lmin_v < lmax_T, "lbg_n0_ho02: need 0 < lmin_v < lmax_T (the integrand is singular at v = 0)");
    LBG_REQUIRE(n_exact >= 0 && (n_exact == 0 || (exact_rows && exact_u)), "lbg_n0_ho02: exact rows given without values");
    LBG_REQUIRE(fold_phi >= 0 && fold_phi <= 2, "lbg_n0_ho02: fold_phi must be 0 (full grid), 1 (mirror pairs) or 2 (mirror pairs, generic kernel)");
    LBG_REQUIRE(c->lane >= 0 && c->lane < LBG_MAX_LANES, "lbg_n0_ho02: context lane %d has no phi-table block (lanes 0..%d)", c->lane, LBG_MAX_LANES - 1);
    c->n0_n_L = n_L;
    c->n0_n_lev = n_lev;
    if (n_lev > 0) deltaT_uK_arcmin = deltaT_lev[0];
    if (n_L == 0) { c->n0_pending = true; return LBG_OK; }
    LBG_CUDA(cudaSetDevice(c->device));
    if (!c->n0_plan) c->n0_plan = new lbg_n0_plan();
    lbg_n0_plan& pl = *c->n0_plan;
    cudaStream_t st = c->stream;

    // ---- plan: grids, row flags, special rows, |u|_j of the special rows -------------------------------------------------
    const bool fold = fold_phi != 0;
    bool plan_fresh = false;
    if (!pl.valid || !pl.key.same(n_l1, l1, w_l1, n_phi, cos_phi, sin_phi, w_phi, lmax_T, lmax_P, lmin_v, fold ? 1 : 0, n_exact, exact_rows, c->n_ell)
        || pl.key.exact_below != exact_below) {
        pl.valid = false;
        plan_fresh = true;
        std::vector<double> ph;   // [6][n_phe]: cos, sin, w (pair sum), mirror cos, mirror sin, w_j / (w_j + w_m)
        int n_phe = 0;
        if (fold) {
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
            for (int j = 0; j < n_phe; ++j) ph[5 * (size_t)n_phe + j] = 1.0;
        }
        LBG_REQUIRE(n_phe + 4 <= LBG_MAX_PHI, "lbg_n0_ho02: %d phi nodes after folding, at most %d fit the constant-memory table", n_phe, LBG_MAX_PHI - 4);
        std::vector<double> phf(3 * (size_t)n_phi);  // full grid for the special rows
        memcpy(&phf[0], cos_phi, (size_t)n_phi * 8);
        memcpy(&phf[n_phi], sin_phi, (size_t)n_phi * 8);
        memcpy(&phf[2 * (size_t)n_phi], w_phi, (size_t)n_phi * 8);
        pl.cos_decreasing = true;
        for (int j = 1; j < n_phe; ++j) pl.cos_decreasing = pl.cos_decreasing && ph[j] < ph[j - 1];
        pl.cw.assign(2 * (size_t)(n_phe + 4), 0.0);
        for (int j = 0; j < n_phe; ++j) { pl.cw[2 * j] = ph[j]; pl.cw[2 * j + 1] = ph[2 * (size_t)n_phe + j]; }

        // row flags; special rows = caller's exact rows, rows below `exact_below` (library splev values), rows on a cut of prep_lmaxmask
        std::vector<int> rowflag(n_l1, 0);
        pl.srow.clear();
        for (int k = 0; k < n_exact; ++k) {
            LBG_REQUIRE(exact_rows[k] >= 0 && exact_rows[k] < n_l1, "lbg_n0_ho02: exact row %d out of range", exact_rows[k]);
            pl.srow.push_back(exact_rows[k]);
        }
        auto have = [&](int r) { for (int q : pl.srow) if (q == r) return true; return false; };
        pl.n_lib_exact = 0;
        for (int r = 0; r < n_l1; ++r)
            if (l1[r] < exact_below && !have(r)) { pl.srow.push_back(r); ++pl.n_lib_exact; }
        pl.active_rows = 0;
        pl.l1_max = 0.0;
        for (int r = 0; r < n_l1; ++r) {
            double l = l1[r];
            LBG_REQUIRE(l > 0.0, "lbg_n0_ho02: l1 nodes must be positive");
            pl.l1_max = std::max(pl.l1_max, l);
            int f = 0;
            bool onT = fabs(l - lmax_T) <= 1e-12 * lmax_T, onP = fabs(l - lmax_P) <= 1e-12 * lmax_P;
            if (l < lmax_T || onT) f |= 1;
            if (l < lmax_P || onP) f |= 4;
            // a row sitting on the LAST KNOT of a table that ends inside the integration range: the reference's interp1d returns the
            // knot value or its fill value 0 per phi node, depending on whether |u|_j = sqrt((l cos)^2 + (l sin)^2) rounds to the knot
            // or one ulp above it (prep_camb.py:88-89) -- the same per-node rounding as the lmax cuts, so the same treatment
            const double l_end = (double)(c->n_ell - 1);
            const bool onEnd = f != 0 && fabs(l - l_end) <= 1e-12 * l_end;
            if ((onT || onP || onEnd) && !have(r)) pl.srow.push_back(r);
            rowflag[r] = f;
            if (f) ++pl.active_rows;
        }
        for (int q : pl.srow) rowflag[q] |= 16;
        pl.n_special = (int)pl.srow.size();
        // |u|_j exactly as cmb/Nkk.py:66-67,79 for every (special row, phi node)
        pl.snu.assign((size_t)pl.n_special * n_phi, 0.0);
        for (int k = 0; k < pl.n_special; ++k)
            for (int j = 0; j < n_phi; ++j) {
                volatile double ux = l1[pl.srow[k]] * cos_phi[j], uy = l1[pl.srow[k]] * sin_phi[j];
                volatile double a2 = ux * ux, b2 = uy * uy;
                volatile double s2 = a2 + b2;
                pl.snu[(size_t)k * n_phi + j] = sqrt(s2);
            }
        pl.n_phe = n_phe;
        // device copies of the plan: one staged block [l1 | w_l1 | ph | phf | snu | rowflag | srow]
        const size_t nd = 2 * (size_t)n_l1 + ph.size() + phf.size() + pl.snu.size();
        const size_t bytes = nd * 8 + ((size_t)n_l1 + pl.n_special) * 4;
        LBG_CUDA(cudaStreamSynchronize(st));       // the staging block may still feed an earlier copy
        if (c->n0_plan_dev.reserve(bytes + 64) || c->pin_n0.reserve(bytes + 64)) return LBG_ENOMEM;
        double* hd = c->pin_n0.as<double>();
        size_t o = 0;
        memcpy(hd + o, l1, (size_t)n_l1 * 8); o += n_l1;
        memcpy(hd + o, w_l1, (size_t)n_l1 * 8); o += n_l1;
        memcpy(hd + o, ph.data(), ph.size() * 8); o += ph.size();
        memcpy(hd + o, phf.data(), phf.size() * 8); o += phf.size();
        if (!pl.snu.empty()) memcpy(hd + o, pl.snu.data(), pl.snu.size() * 8);
        o += pl.snu.size();
        int* hi = reinterpret_cast<int*>(hd + o);
        memcpy(hi, rowflag.data(), (size_t)n_l1 * 4);
        if (pl.n_special) memcpy(hi + n_l1, pl.srow.data(), (size_t)pl.n_special * 4);
        LBG_CUDA(cudaMemcpyAsync(c->n0_plan_dev.p, hd, bytes, cudaMemcpyHostToDevice, st));
        LBG_CUDA(cudaStreamSynchronize(st));       // pin_n0 is reused for the per-call block below
        N0Key& k = pl.key;
        k.n_l1 = n_l1; k.n_phi = n_phi; k.fold = fold ? 1 : 0; k.n_exact = n_exact; k.n_ell = c->n_ell;
        k.lmax_T = lmax_T; k.lmax_P = lmax_P; k.lmin_v = lmin_v; k.exact_below = exact_below;
        k.l1.assign(l1, l1 + n_l1); k.w_l1.assign(w_l1, w_l1 + n_l1);
        k.cphi.assign(cos_phi, cos_phi + n_phi); k.sphi.assign(sin_phi, sin_phi + n_phi); k.wphi.assign(w_phi, w_phi + n_phi);
        k.exact_rows.assign(exact_rows, exact_rows + n_exact);
        c->cw_lane_valid = false;
        pl.valid = true;
    }
    const int n_phe = pl.n_phe, n_special = pl.n_special;
    const double* d_l1 = c->n0_plan_dev.as<double>();
    const double* d_wl1 = d_l1 + n_l1;
    const double* d_ph = d_wl1 + n_l1;
    const double* d_phf = d_ph + 6 * (size_t)n_phe;
    const double* d_snu = d_phf + 3 * (size_t)n_phi;
    const int* d_rowflag = reinterpret_cast<const int*>(d_snu + (size_t)n_special * n_phi);
    const int* d_srow = d_rowflag + n_l1;
    bool cw_mine;
    {
        std::lock_guard<std::mutex> g(g_cw_mu);
        lbg_ctx*& owner = g_cw_owner[c->device & 63][c->lane];
        cw_mine = owner == c;
        owner = c;
    }
    if (!c->cw_lane_valid || !cw_mine) {
        LBG_CUDA(cudaMemcpyToSymbolAsync(c_cw, pl.cw.data(), pl.cw.size() * 8, (size_t)c->lane * LBG_MAX_PHI * sizeof(double2),
                                         cudaMemcpyHostToDevice, st));
        LBG_CUDA(cudaStreamSynchronize(st));
        c->cw_lane_valid = true;
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
    P.n_rb = (n_l1 + 127) / 128;
    P.lane = c->lane;
    P.nT = (dT / Tcmb_K) * (dT / Tcmb_K);
    P.nP = (dP / Tcmb_K) * (dP / Tcmb_K);
    P.a = tb * tb / (8. * log(2.));
    P.lmax_T = lmax_T; P.lmax_P = lmax_P; P.lmin_v = lmin_v;
    P.ebfac = iterative ? 2.5 : 1.0;                              // cmb/Nkk.py:141-143
    P.fold = fold ? 1 : 0;
    // every |v| an unmasked node can have lies below vmax: the polynomial table covers [0, vmax]
    double Lmax = 0.0;
    for (int i = 0; i < n_L; ++i) { LBG_REQUIRE(L[i] == L[i], "lbg_n0_ho02: L[%d] is NaN", i); Lmax = std::max(Lmax, fabs(L[i])); }
    const double vmax = std::min(lmax_P, Lmax + pl.l1_max);
    LBG_REQUIRE(vmax + 3.0 < (double)LBG_PAIR_STRIDE, "lbg_n0_ho02: |v| reaches %.0f; the per-call spline table holds %d knot intervals "
                "(lower lmax_P or L)", vmax, LBG_PAIR_STRIDE - 3);
    const bool short_table = (double)(c->n_ell - 1) < vmax;     // the table ends inside the reachable |v| range
    P.n_int = short_table ? (int)vmax + 2 : std::min(c->n_ell - 1, LBG_PAIR_STRIDE);
    {
        const double thr[4] = {lmin_v, lmax_T, lmax_P, (double)(c->n_ell - 1)};
        for (int k = 0; k < 4; ++k) {
            double lo = thr[k] * thr[k] * (1.0 - 4e-6), hi = thr[k] * thr[k] * (1.0 + 4e-6);
            unsigned long long blo, bhi;
            memcpy(&blo, &lo, 8); memcpy(&bhi, &hi, 8);
            P.near_lo[k] = (unsigned)(blo >> 32);
            P.near_span[k] = (unsigned)(bhi >> 32) - P.near_lo[k];
        }
        if (!short_table) { P.near_lo[3] = 0xffffffffu; P.near_span[3] = 0u; }    // (hi - lo) <= span never holds for a finite |v|^2
    }
    const long long npairs = (long long)n_L * n_phe, npairs_full = (long long)n_L * n_phi;
    LBG_REQUIRE(npairs_full < (1LL << 30), "lbg_n0_ho02: n_L * n_phi too large");
    LBG_REQUIRE((long long)n_L * P.n_rb < (1LL << 31), "lbg_n0_ho02: n_L * row blocks exceeds the grid limit");

    // ---- which integrand kernel ---------------------------------------------------------------------------------------------
    // folded-noise bound of the per-interval Chebyshev cubic (see the table's header)
    const double beta = P.a * (2.0 * vmax + 1.0);
    const double bound = (beta * beta * beta * beta + 12.0 * P.a * beta * beta + 12.0 * P.a * P.a) / 3072.0;
    const bool expn = bound > 2e-12;
    const bool rows5 = fold_phi == 1 && pl.cos_decreasing && !short_table && !expn &&
                       P.near_lo[0] + P.near_span[0] < P.near_lo[1] && P.near_lo[1] <= P.near_lo[2];
    c->n0_path = rows5 ? 5 : 3;
    ChebFit fit;
    {   // monomial <- samples matrix of the four Chebyshev nodes on [0, 1] (Gauss-Jordan in long double)
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

    // ---- per-call staging block: [L | u values of the special rows] in pinned memory, one copy -----------------------------------
    const size_t su_count = (size_t)n_special * n_phi * 8;
    const size_t part_main = (size_t)6 * n_L * P.n_rb, part_special = (size_t)6 * npairs_full;
    const size_t nlv = (size_t)std::max(1, n_lev);
    if (c->n0_L.reserve((size_t)n_L * 8 + su_count * 8 + 64) || c->pin_n0.reserve((size_t)n_L * 8 + su_count * 8 + 64) ||
        c->n0_utab.reserve(nlv * n_l1 * sizeof(URow)) || c->n0_partial.reserve(nlv * (part_main + part_special) * 8) ||
        c->n0_out.reserve(nlv * 7 * n_L * 8) || c->pin_n0_out.reserve(nlv * 7 * n_L * 8) ||
        c->cl_poly.reserve((size_t)LBG_PAIR_STRIDE * 8 * 36))
        return LBG_ENOMEM;
    double* hL = c->pin_n0.as<double>();
    memcpy(hL, L, (size_t)n_L * 8);
    double* hsu = hL + n_L;
    if (n_exact) memcpy(hsu, exact_u, (size_t)n_exact * n_phi * 64);
    if (pl.n_lib_exact) {
        // u-side values of the rows below `exact_below` by the reference's own spline routine (deboor_cubic), cached per table upload
        const size_t cnt = (size_t)pl.n_lib_exact * n_phi * 8;
        if (plan_fresh || c->n0_exact_gen != c->cl_gen || c->n0_exact_vals.size() != cnt) {
            if (int rc = ensure_scipy_bspline(c)) return rc;
            c->n0_exact_vals.assign(cnt, 0.0);
            const int nk = c->n_ell;
            const double xmax = (double)(nk - 1);
            for (int k = 0; k < pl.n_lib_exact; ++k)
                for (int j = 0; j < n_phi; ++j) {
                    const double x = pl.snu[(size_t)(n_exact + k) * n_phi + j];
                    for (int s8 = 0; s8 < 8; ++s8) {
                        double v = 0.0;                                   // bounds_error=False, fill_value=0.0
                        if (x >= 0.0 && x <= xmax) v = scipy_bspline_eval(c->cl_bspl_t.data(), c->cl_bspl_c.data() + (size_t)s8 * nk, nk, x);
                        c->n0_exact_vals[((size_t)k * n_phi + j) * 8 + s8] = v;
                    }
                }
            c->n0_exact_gen = c->cl_gen;
        }
        memcpy(hsu + (size_t)n_exact * n_phi * 8, c->n0_exact_vals.data(), cnt * 8);
    }
    const int n_given = n_exact + pl.n_lib_exact;                // special rows whose u values arrive from the host
    double* d_L = c->n0_L.as<double>();
    double* d_su = d_L + n_L;
    LBG_CUDA(cudaMemcpyAsync(d_L, hL, ((size_t)n_L + (size_t)n_given * n_phi * 8) * 8, cudaMemcpyHostToDevice, st));
    double* d_part = c->n0_partial.as<double>();
    double* d_spart = d_part + part_main;

    if (n_lev > 0) {
        // ---- several noise levels of one beam: per-level u rows / special rows / reductions, the integrand in groups of <= 3 levels ----
        LBG_REQUIRE(rows5 && terms_mask == (LBG_TT | LBG_TE | LBG_EE | LBG_EB),
                    "lbg_n0_ho02_levels: needs the four estimators TT,TE,EE,EB, a folded phi grid with decreasing cos, tables that cover every "
                    "reachable |v| and a beam inside the folded-noise bound (use lbg_n0_ho02 per level otherwise)");
        c->n0_path = 6;
        c->last_launches = 0;
        cudaEventRecord(c->ev0, st);
        if (n_special > n_given) {
            int cnt = (n_special - n_given) * n_phi;
            n0_special_u_kernel<<<(cnt + 127) / 128, 128, 0, st>>>(c->n_ell, cnt, d_snu + (size_t)n_given * n_phi,
                                                                   c->cl_rows.as<ClRow>(), d_su + (size_t)n_given * n_phi * 8);
            ++c->last_launches;
        }
        n0_poly_table_kernel<<<(P.n_int + 127) / 128, 128, 0, st>>>(P, fit, 2, c->cl_soa.as<double>(), c->cl_poly.as<double4>());
        ++c->last_launches;
        std::vector<N0Params> PL(n_lev, P);
        for (int v = 0; v < n_lev; ++v) {
            double dTv = deltaT_lev[v] * pow(10., -6.);
            dTv *= arcmin2rad;
            const double dPv = sqrt(2.) * dTv;
            PL[v].nT = (dTv / Tcmb_K) * (dTv / Tcmb_K);
            PL[v].nP = (dPv / Tcmb_K) * (dPv / Tcmb_K);
            PL[v].ebfac = (iterative_lev && iterative_lev[v]) ? 2.5 : 1.0;
            n0_uside_kernel<<<(n_l1 + 127) / 128, 128, 0, st>>>(PL[v], d_l1, d_wl1, d_rowflag, c->cl_rows.as<ClRow>(),
                                                                 c->n0_utab.as<URow>() + (size_t)v * n_l1);
            ++c->last_launches;
        }
        cudaEventRecord(c->ev2, st);
        const unsigned grid = (unsigned)((long long)n_L * P.n_rb);
        for (int v0 = 0; v0 < n_lev; v0 += LBG_MAX_NLEV) {
            const int nl = std::min(LBG_MAX_NLEV, n_lev - v0);
            LevelParams LP;
            for (int q = 0; q < LBG_MAX_NLEV; ++q) { const N0Params& S = PL[v0 + std::min(q, nl - 1)]; LP.lv[q].nT = S.nT; LP.lv[q].nP = S.nP; }
            const URow* ut = c->n0_utab.as<URow>() + (size_t)v0 * n_l1;
            double* part = d_part + (size_t)v0 * part_main;
#define LBG_R5M(NL, UB) n0_rows5m_kernel<NL, UB><<<grid, 128, 0, st>>>(P, LP, d_L, ut, c->cl_poly.as<double4>(), c->cl_soa.as<double>(), d_ph, part)
            if (c->ub_zero) { if (nl == 3) LBG_R5M(3, false); else if (nl == 2) LBG_R5M(2, false); else LBG_R5M(1, false); }
            else            { if (nl == 3) LBG_R5M(3, true);  else if (nl == 2) LBG_R5M(2, true);  else LBG_R5M(1, true); }
#undef LBG_R5M
            ++c->last_launches;
        }
        cudaEventRecord(c->ev3, st);
        double* d_spart_all = d_part + (size_t)n_lev * part_main;
        for (int v = 0; v < n_lev; ++v) {
            double* sp = d_spart_all + (size_t)v * part_special;
            if (n_special) {
                n0_special_kernel<<<(int)((npairs_full + 127) / 128), 128, 0, st>>>(PL[v], n_special, n_phi, d_srow, d_L, d_phf,
                                                                                    c->n0_utab.as<URow>() + (size_t)v * n_l1, d_su, d_snu,
                                                                                    c->cl_rows.as<ClRow>(), sp);
                ++c->last_launches;
            }
            n0_rows_reduce_kernel<<<n_L, 128, 0, st>>>(PL[v], d_L, d_part + (size_t)v * part_main, n_special ? sp : nullptr, n_phi,
                                                       c->n0_out.as<double>() + (size_t)v * 7 * n_L);
            ++c->last_launches;
        }
        cudaEventRecord(c->ev1, st);
        LBG_CUDA(cudaMemcpyAsync(c->pin_n0_out.p, c->n0_out.p, (size_t)n_lev * 7 * n_L * 8, cudaMemcpyDeviceToHost, st));
        LBG_CUDA(cudaGetLastError());
        c->n0_nodes_walked = npairs * (pl.active_rows - n_special) + npairs_full * n_special;
        c->n0_pending = true;
        return LBG_OK;
    }

    c->last_launches = 0;
    cudaEventRecord(c->ev0, st);
    n0_uside_kernel<<<(n_l1 + 127) / 128, 128, 0, st>>>(P, d_l1, d_wl1, d_rowflag, c->cl_rows.as<ClRow>(), c->n0_utab.as<URow>());
    ++c->last_launches;
    if (n_special > n_given) {   // device-evaluated splines at |u|_j for the rows on a cut
        int cnt = (n_special - n_given) * n_phi;
        n0_special_u_kernel<<<(cnt + 127) / 128, 128, 0, st>>>(c->n_ell, cnt, d_snu + (size_t)n_given * n_phi,
                                                               c->cl_rows.as<ClRow>(), d_su + (size_t)n_given * n_phi * 8);
        ++c->last_launches;
    }
    n0_poly_table_kernel<<<(P.n_int + 127) / 128, 128, 0, st>>>(P, fit, expn ? 0 : 1, c->cl_soa.as<double>(), c->cl_poly.as<double4>());
    ++c->last_launches;
    cudaEventRecord(c->ev2, st);
    const unsigned grid = (unsigned)((long long)n_L * P.n_rb);
    const bool full = terms_mask == (LBG_TT | LBG_TE | LBG_EE | LBG_EB);
#define LBG_R5(UB, FULL) n0_rows5_kernel<UB, FULL><<<grid, 128, 0, st>>>(P, d_L, c->n0_utab.as<URow>(), c->cl_poly.as<double4>(), \
                             c->cl_soa.as<double>(), d_ph, d_part)
#define LBG_R3(EXPN) n0_rows3_kernel<EXPN><<<grid, 128, 0, st>>>(P, d_L, c->n0_utab.as<URow>(), c->cl_poly.as<double4>(), \
                             c->cl_soa.as<double>(), d_ph, d_part)
    if (rows5) {
        if (c->ub_zero) { if (full) LBG_R5(false, true); else LBG_R5(false, false); }
        else            { if (full) LBG_R5(true, true);  else LBG_R5(true, false); }
    } else {
        if (expn) LBG_R3(true); else LBG_R3(false);
    }
#undef LBG_R5
#undef LBG_R3
    ++c->last_launches;
    cudaEventRecord(c->ev3, st);
    if (n_special) {
        n0_special_kernel<<<(int)((npairs_full + 127) / 128), 128, 0, st>>>(P, n_special, n_phi, d_srow, d_L, d_phf, c->n0_utab.as<URow>(),
                                                                            d_su, d_snu, c->cl_rows.as<ClRow>(), d_spart);
        ++c->last_launches;
    }
    n0_rows_reduce_kernel<<<n_L, 128, 0, st>>>(P, d_L, d_part, n_special ? d_spart : nullptr, n_phi, c->n0_out.as<double>());
    ++c->last_launches;
    cudaEventRecord(c->ev1, st);
    LBG_CUDA(cudaMemcpyAsync(c->pin_n0_out.p, c->n0_out.p, (size_t)7 * n_L * 8, cudaMemcpyDeviceToHost, st));
    LBG_CUDA(cudaGetLastError());
    c->n0_nodes_walked = npairs * (pl.active_rows - n_special) + npairs_full * n_special;
    c->n0_pending = true;
    return LBG_OK;
}

static int n0_end_locked(lbg_ctx* c, int n_L, double* out_nkk, double* out_terms, long long* nodes_executed, long long* nodes_unmasked,
                         int n_lev = 0) {
    if (!c->n0_pending) { set_error("lbg_n0_ho02_end: no pending call on this context"); return LBG_ESTATE; }
    c->n0_pending = false;
    if (n_L != c->n0_n_L || n_lev != c->n0_n_lev) {   // the pending call is dropped; wait for it so that its buffers can be reused
        cudaSetDevice(c->device);
        cudaStreamSynchronize(c->stream);
        set_error("lbg_n0_ho02_end: n_L=%d, n_lev=%d but the pending call has %d multipoles, %d levels (call dropped)", n_L, n_lev, c->n0_n_L, c->n0_n_lev);
        return LBG_EINVAL;
    }
    if (n_L == 0) { if (nodes_executed) *nodes_executed = 0; if (nodes_unmasked) *nodes_unmasked = 0; return LBG_OK; }
    LBG_REQUIRE(out_nkk, "lbg_n0_ho02_end: out_nkk is NULL");
    LBG_CUDA(cudaSetDevice(c->device));
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    if (e != cudaSuccess) { set_error("kernel execution failed: %s", cudaGetErrorString(e)); return LBG_ECUDA; }
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, c->ev0, c->ev1) == cudaSuccess) { c->last_ms = ms; c->total_ms += ms; c->total_launches += c->last_launches; }
    c->last_main_ms = cudaEventElapsedTime(&ms, c->ev2, c->ev3) == cudaSuccess ? ms : c->last_ms;
    const double* host = c->pin_n0_out.as<double>();
    for (int v = 0; v < std::max(1, n_lev); ++v) {      // per level: [N0 | 5 estimator integrals | unmasked count], each n_L long
        const double* hv = host + (size_t)v * 7 * n_L;
        memcpy(out_nkk + (size_t)v * n_L, hv, (size_t)n_L * 8);
        if (out_terms) memcpy(out_terms + (size_t)v * 5 * n_L, hv + n_L, (size_t)5 * n_L * 8);
    }
    if (nodes_executed) *nodes_executed = c->n0_nodes_walked;
    if (nodes_unmasked) {
        double s = 0.0;
        for (int i = 0; i < n_L; ++i) s += host[(size_t)6 * n_L + i];
        *nodes_unmasked = (long long)(s + 0.5);
    }
    return LBG_OK;
}

extern "C" int lbg_n0_ho02(lbg_ctx* c, int n_L, const double* L, unsigned terms_mask,
                           double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                           int n_l1, const double* l1, const double* w_l1,
                           int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                           double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                           int n_exact, const int* exact_rows, const double* exact_u,
                           double* out_nkk, double* out_terms, long long* nodes_executed) {
    LBG_REQUIRE(c && out_nkk, "lbg_n0_ho02: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    int rc = n0_begin_locked(c, n_L, L, terms_mask, thetab_arcmin, deltaT_uK_arcmin, Tcmb_K, iterative, n_l1, l1, w_l1, n_phi, cos_phi, sin_phi,
                             w_phi, lmax_T, lmax_P, lmin_v, fold_phi, n_exact, exact_rows, exact_u, c->n0_exact_below);
    if (rc) return rc;
    return n0_end_locked(c, n_L, out_nkk, out_terms, nodes_executed, nullptr);
}

extern "C" int lbg_n0_ho02_begin(lbg_ctx* c, int n_L, const double* L, unsigned terms_mask,
                                 double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                                 int n_l1, const double* l1, const double* w_l1,
                                 int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                                 double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                                 int n_exact, const int* exact_rows, const double* exact_u) {
    LBG_REQUIRE(c, "lbg_n0_ho02_begin: ctx is NULL");
    std::lock_guard<std::mutex> g(c->mu);
    return n0_begin_locked(c, n_L, L, terms_mask, thetab_arcmin, deltaT_uK_arcmin, Tcmb_K, iterative, n_l1, l1, w_l1, n_phi, cos_phi, sin_phi,
                           w_phi, lmax_T, lmax_P, lmin_v, fold_phi, n_exact, exact_rows, exact_u, c->n0_exact_below);
}

extern "C" int lbg_n0_ho02_end(lbg_ctx* c, int n_L, double* out_nkk, double* out_terms, long long* nodes_executed, long long* nodes_unmasked) {
    LBG_REQUIRE(c, "lbg_n0_ho02_end: ctx is NULL");
    std::lock_guard<std::mutex> g(c->mu);
    return n0_end_locked(c, n_L, out_nkk, out_terms, nodes_executed, nodes_unmasked);
}

extern "C" int lbg_n0_set_exact_below(lbg_ctx* c, double l_below) {
    LBG_REQUIRE(c, "lbg_n0_set_exact_below: ctx is NULL");
    std::lock_guard<std::mutex> g(c->mu);
    c->n0_exact_below = l_below;
    return LBG_OK;
}

extern "C" int lbg_n0_last_path(lbg_ctx* c, int* path) {
    LBG_REQUIRE(c && path, "lbg_n0_last_path: NULL argument");
    *path = c->n0_path;
    return LBG_OK;
}

void lbg_n0_release(lbg_ctx* c) { n0_plan_free(c); }

// ---- several noise levels of one beam in one pass (BASELINE configs[4]: beam x noise sweeps; snr.py:108-124 loops Nkk over experiments) ----
static int check_levels(int n_lev, const double* deltaT, const char* who) {
    LBG_REQUIRE(n_lev >= 1 && n_lev <= 4096 && deltaT, "%s: need 1 <= n_lev <= 4096 and the noise levels", who);
    for (int v = 0; v < n_lev; ++v) LBG_REQUIRE(deltaT[v] >= 0.0 && deltaT[v] == deltaT[v], "%s: deltaT[%d] is negative or NaN", who, v);
    return LBG_OK;
}

extern "C" int lbg_n0_ho02_levels_begin(lbg_ctx* c, int n_L, const double* L, double thetab_arcmin, int n_lev, const double* deltaT_uK_arcmin,
                                        const int* iterative, double Tcmb_K,
                                        int n_l1, const double* l1, const double* w_l1,
                                        int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                                        double lmax_T, double lmax_P, double lmin_v) {
    LBG_REQUIRE(c, "lbg_n0_ho02_levels_begin: ctx is NULL");
    if (int rc = check_levels(n_lev, deltaT_uK_arcmin, "lbg_n0_ho02_levels")) return rc;
    std::lock_guard<std::mutex> g(c->mu);
    return n0_begin_locked(c, n_L, L, LBG_TT | LBG_TE | LBG_EE | LBG_EB, thetab_arcmin, 0.0, Tcmb_K, 0, n_l1, l1, w_l1, n_phi, cos_phi, sin_phi,
                           w_phi, lmax_T, lmax_P, lmin_v, 1, 0, nullptr, nullptr, c->n0_exact_below, n_lev, deltaT_uK_arcmin, iterative);
}

extern "C" int lbg_n0_ho02_levels_end(lbg_ctx* c, int n_lev, int n_L, double* out_nkk, double* out_terms) {
    LBG_REQUIRE(c && n_lev >= 1, "lbg_n0_ho02_levels_end: bad argument");
    std::lock_guard<std::mutex> g(c->mu);
    return n0_end_locked(c, n_L, out_nkk, out_terms, nullptr, nullptr, n_lev);
}

extern "C" int lbg_n0_ho02_levels(lbg_ctx* c, int n_L, const double* L, double thetab_arcmin, int n_lev, const double* deltaT_uK_arcmin,
                                  const int* iterative, double Tcmb_K,
                                  int n_l1, const double* l1, const double* w_l1,
                                  int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                                  double lmax_T, double lmax_P, double lmin_v, double* out_nkk, double* out_terms) {
    int rc = lbg_n0_ho02_levels_begin(c, n_L, L, thetab_arcmin, n_lev, deltaT_uK_arcmin, iterative, Tcmb_K, n_l1, l1, w_l1, n_phi, cos_phi, sin_phi,
                                      w_phi, lmax_T, lmax_P, lmin_v);
    if (rc) return rc;
    return lbg_n0_ho02_levels_end(c, n_lev, n_L, out_nkk, out_terms);
}
