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
    double flags;                 // bit0 T-terms active, bit1 T boundary row, bit2 P active, bit3 P boundary row
    double pad0, pad1;
};

struct N0Params {
    int n_L, n_phi, n_l1, n_chunk, rows_per_chunk, n_ell;
    unsigned terms;
    double nT, nP, a;             // N_TT(l) = nT exp(a l (l+1)), N_EE = N_BB = nP exp(.)  (cmb/DetectorNoise.py:22-26)
    double lmax_T, lmax_P, lmin_v;
    double ebfac;
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

__global__ void __launch_bounds__(128)
n0_kernel(N0Params P, const double* __restrict__ Ls, const double* __restrict__ phi /*[5][n_phi]: cos,sin,w,ratioT,ratioP*/,
          const URow* __restrict__ utab, const ClRow* __restrict__ rows, double* __restrict__ partial) {
    const int npairs = P.n_L * P.n_phi;
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = p < npairs;
    const int pp = valid ? p : npairs - 1;
    const int iL = pp / P.n_phi, j = pp - iL * P.n_phi;
    const double L = Ls[iL], c = phi[j], s = phi[P.n_phi + j];
    const int r0 = blockIdx.y * P.rows_per_chunk;
    const int r1 = min(P.n_l1, r0 + P.rows_per_chunk);
    const double xmax = (double)(P.n_ell - 1);
    const unsigned terms = P.terms;
    double aTT = 0.0, aTE = 0.0, aEE = 0.0, aEB = 0.0, aBB = 0.0;

    for (int r = r0; r < r1; ++r) {
        const double4* up = reinterpret_cast<const double4*>(utab + r);
        const double4 q3 = ld4(up + 3);              // rB1h, flags, pad, pad
        const int fl = (int)q3.y;
        if (!(fl & 5)) continue;                       // block-uniform: row masked for every estimator
        const double4 q0 = ld4(up);                  // l, wl, uT, uX
        const double l = q0.x;
        // geometry exactly as cmb/Nkk.py:66-72,80 (no FMA contraction: |v| decides the masks)
        const double ux = __dmul_rn(l, c), uy = __dmul_rn(l, s);
        const double vx = __dsub_rn(L, ux);
        const double nv2 = __dadd_rn(__dmul_rn(vx, vx), __dmul_rn(uy, uy));
        const double nv = __dsqrt_rn(nv2);
        const bool vok = !(nv < P.lmin_v);
        const bool okT = (fl & 1) && vok && !(nv > P.lmax_T);
        const bool okP = (fl & 4) && vok && !(nv > P.lmax_P);
        if (!(okT || okP)) continue;
        double mT = okT ? 1.0 : 0.0, mP = okP ? 1.0 : 0.0;
        if (fl & 2) mT *= phi[3 * P.n_phi + j];        // boundary row: |u| > lmax_T decided per phi node
        if (fl & 8) mP *= phi[4 * P.n_phi + j];
        const double4 q1 = ld4(up + 1);              // uE, uB, T1, E1
        const double4 q2 = ld4(up + 2);              // X1, rT1h, rE1h, rE1
        const double wl = q0.y, uT = q0.z, uX = q0.w, uE = q1.x, uB = q1.y, T1 = q1.z, E1 = q1.w;
        const double X1 = q2.x, rT1h = q2.y, rE1h = q2.z, rE1 = q2.w, rB1h = q3.x;

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
        if (nv > xmax) { w0 = w1 = w2 = w3 = 0.0; }
        const double4* rp = reinterpret_cast<const double4*>(rows + i);
        double4 A = ld4(rp), B = ld4(rp + 1);
        double vT = w0 * A.x, vX = w0 * A.y, vE = w0 * A.z, vB = w0 * A.w;
        double T2 = w0 * B.x, X2 = w0 * B.y, E2 = w0 * B.z, B2 = w0 * B.w;
        A = ld4(rp + 2); B = ld4(rp + 3);
        vT = fma(w1, A.x, vT); vX = fma(w1, A.y, vX); vE = fma(w1, A.z, vE); vB = fma(w1, A.w, vB);
        T2 = fma(w1, B.x, T2); X2 = fma(w1, B.y, X2); E2 = fma(w1, B.z, E2); B2 = fma(w1, B.w, B2);
        A = ld4(rp + 4); B = ld4(rp + 5);
        vT = fma(w2, A.x, vT); vX = fma(w2, A.y, vX); vE = fma(w2, A.z, vE); vB = fma(w2, A.w, vB);
        T2 = fma(w2, B.x, T2); X2 = fma(w2, B.y, X2); E2 = fma(w2, B.z, E2); B2 = fma(w2, B.w, B2);
        A = ld4(rp + 6); B = ld4(rp + 7);
        vT = fma(w3, A.x, vT); vX = fma(w3, A.y, vX); vE = fma(w3, A.z, vE); vB = fma(w3, A.w, vB);
        T2 = fma(w3, B.x, T2); X2 = fma(w3, B.y, X2); E2 = fma(w3, B.z, E2); B2 = fma(w3, B.w, B2);
        // detector noise at |v| (cmb/DetectorNoise.py:22-26)
        const double en = exp(nv * (nv + 1.0) * P.a);
        T2 = fma(P.nT, en, T2);
        E2 = fma(P.nP, en, E2);
        B2 = fma(P.nP, en, B2);

        if (okT) {
            const double wT = wl * mT;
            if (terms & LBG_TT) {                      // cmb/Nkk.py:93-99
                const double fa = uT * dLu + vT * dLv;
                aTT = fma(wT, fa * fa * rT1h / T2, aTT);
            }
            if (terms & LBG_TE) {                      // cmb/Nkk.py:101-125
                const double fa_uv = uX * c2 * dLu + vX * dLv;
                const double fa_vu = uX * dLu + vX * c2 * dLv;
                const double xx = X1 * X2;
                const double num = E1 * T2 * fa_uv - xx * fa_vu;
                const double den = T1 * E2 * E1 * T2 - xx * xx;
                aTE = fma(wT, fa_uv * num / den, aTE);
            }
        }
        if (okP) {
            const double wP = wl * mP;
            if (terms & LBG_EE) {                      // cmb/Nkk.py:127-133
                const double fa = (uE * dLu + vE * dLv) * c2;
                aEE = fma(wP, fa * fa * rE1h / E2, aEE);
            }
            if (terms & LBG_EB) {                      // cmb/Nkk.py:135-146
                const double fa = uE * dLu - vB * dLv;
                aEB = fma(wP, fa * fa * s2sq * rE1 / B2, aEB);
            }
            if (terms & LBG_BB) {                      // cmb/Nkk.py:148-154
                const double fa = (uB * dLu + vB * dLv) * c2;
                aBB = fma(wP, fa * fa * rB1h / B2, aBB);
            }
        }
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

// One block per L: fixed-order reduction of the partial sums, then cmb/Nkk.py:169-172.
__global__ void __launch_bounds__(128)
n0_reduce_kernel(N0Params P, const double* __restrict__ Ls, const double* __restrict__ partial,
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
                           double* out_nkk, double* out_terms, long long* nodes_executed) {
    LBG_REQUIRE(c && L && l1 && w_l1 && cos_phi && sin_phi && w_phi && out_nkk, "lbg_n0_ho02: NULL argument");
    LBG_REQUIRE(c->n_ell > 0, "lbg_n0_ho02: call lbg_upload_cl_tables first");
    if (c->n_ell <= 0) return LBG_ESTATE;
    LBG_REQUIRE(n_L >= 0 && n_l1 > 0 && n_phi > 0, "lbg_n0_ho02: bad sizes n_L=%d n_l1=%d n_phi=%d", n_L, n_l1, n_phi);
    LBG_REQUIRE((terms_mask & ~(LBG_TT | LBG_TE | LBG_EE | LBG_EB | LBG_BB)) == 0 && terms_mask != 0,
                "lbg_n0_ho02: unsupported terms mask 0x%x", terms_mask);
    LBG_REQUIRE(Tcmb_K > 0 && lmax_T <= lmax_P, "lbg_n0_ho02: need Tcmb > 0 and lmax_T <= lmax_P");
    if (n_L == 0) { if (nodes_executed) *nodes_executed = 0; return LBG_OK; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));

    // ---- phi nodes (optionally folded onto one node per mirror pair) -------------------------------------
    std::vector<int> rep, mir;                       // representative node, its mirror (== rep if self-mirrored / unfolded)
    if (fold_phi) {
        for (int j = 0, m = n_phi - 1; j <= m; ++j, --m) {
            bool ok = fabs(cos_phi[j] - cos_phi[m]) < 1e-9 && fabs(sin_phi[j] + sin_phi[m]) < 1e-9;
            LBG_REQUIRE(ok, "lbg_n0_ho02: fold_phi requested but nodes %d and %d are not mirror images", j, m);
            rep.push_back(j); mir.push_back(m);
        }
    } else {
        for (int j = 0; j < n_phi; ++j) { rep.push_back(j); mir.push_back(j); }
    }
    const int n_phe = (int)rep.size();

    // ---- row flags; |u| masks of prep_lmaxmask (cmb/Nkk.py:17,20) on rows that sit on a cut ---------------
    std::vector<int> rowflag(n_l1, 0);
    int bT = -1, bP = -1;
    long long active_rows = 0;
    for (int r = 0; r < n_l1; ++r) {
        double l = l1[r];
        LBG_REQUIRE(l > 0.0, "lbg_n0_ho02: l1 nodes must be positive");
        int f = 0;
        if (l < lmax_T * (1.0 - 1e-12)) f |= 1; else if (l <= lmax_T * (1.0 + 1e-12)) { f |= 3; LBG_REQUIRE(bT < 0, "two l1 nodes on lmax_T"); bT = r; }
        if (l < lmax_P * (1.0 - 1e-12)) f |= 4; else if (l <= lmax_P * (1.0 + 1e-12)) { f |= 12; LBG_REQUIRE(bP < 0, "two l1 nodes on lmax_P"); bP = r; }
        rowflag[r] = f;
        if (f) ++active_rows;
    }
    std::vector<double> ph(5 * (size_t)n_phe, 1.0);
    auto umask = [&](int row, int node, double thr) -> double {   // 1 if the reference keeps |u| at this node
        volatile double ux = l1[row] * cos_phi[node], uy = l1[row] * sin_phi[node];
        volatile double a2 = ux * ux, b2 = uy * uy;
        volatile double s2 = a2 + b2;
        double nu = sqrt(s2);                                     // cmb/Nkk.py:79
        return nu > thr ? 0.0 : 1.0;
    };
    for (int e = 0; e < n_phe; ++e) {
        int j = rep[e], m = mir[e];
        double w = (m != j) ? w_phi[j] + w_phi[m] : w_phi[j];
        ph[e] = cos_phi[j]; ph[n_phe + e] = sin_phi[j]; ph[2 * (size_t)n_phe + e] = w;
        for (int k = 0; k < 2; ++k) {
            int row = k ? bP : bT; double thr = k ? lmax_P : lmax_T;
            double ratio = 1.0;
            if (row >= 0) {
                double kept = w_phi[j] * umask(row, j, thr) + (m != j ? w_phi[m] * umask(row, m, thr) : 0.0);
                ratio = (w != 0.0) ? kept / w : 0.0;
            }
            ph[(3 + k) * (size_t)n_phe + e] = ratio;
        }
    }

    // ---- noise constants, cmb/DetectorNoise.py:13-26 ------------------------------------------------------
    const double arcmin2rad = M_PI / (60. * 180.);
    double dT = deltaT_uK_arcmin * pow(10., -6.);
    dT *= arcmin2rad;
    const double tb = thetab_arcmin * arcmin2rad;
    const double dP = sqrt(2.) * dT;

    N0Params P;
    P.n_L = n_L; P.n_phi = n_phe; P.n_l1 = n_l1; P.n_ell = c->n_ell; P.terms = terms_mask;
    P.nT = (dT / Tcmb_K) * (dT / Tcmb_K);
    P.nP = (dP / Tcmb_K) * (dP / Tcmb_K);
    P.a = tb * tb / (8. * log(2.));
    P.lmax_T = lmax_T; P.lmax_P = lmax_P; P.lmin_v = lmin_v;
    P.ebfac = iterative ? 2.5 : 1.0;                              // cmb/Nkk.py:141-143
    const long long npairs = (long long)n_L * n_phe;
    LBG_REQUIRE(npairs < (1LL << 30), "lbg_n0_ho02: n_L * n_phi too large");
    const int bx = (int)((npairs + 127) / 128);
    int n_chunk = (c->sm_count * 32 + bx - 1) / bx;
    n_chunk = std::max(1, std::min(n_chunk, std::max(1, n_l1 / 128)));
    P.rows_per_chunk = (n_l1 + n_chunk - 1) / n_chunk;
    n_chunk = (n_l1 + P.rows_per_chunk - 1) / P.rows_per_chunk;
    P.n_chunk = n_chunk;

    if (c->n0_L.reserve((size_t)n_L * 8) || c->s0.reserve((size_t)n_l1 * 8) || c->s1.reserve((size_t)n_l1 * 8) ||
        c->n0_rowflag.reserve((size_t)n_l1 * 4) || c->n0_phi.reserve(ph.size() * 8) ||
        c->n0_utab.reserve((size_t)n_l1 * sizeof(URow)) ||
        c->n0_partial.reserve((size_t)5 * n_chunk * npairs * 8) || c->n0_out.reserve((size_t)6 * n_L * 8))
        return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->n0_L.p, L, (size_t)n_L * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, l1, (size_t)n_l1 * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, w_l1, (size_t)n_l1 * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->n0_rowflag.p, rowflag.data(), (size_t)n_l1 * 4, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->n0_phi.p, ph.data(), ph.size() * 8, cudaMemcpyHostToDevice, st));

    KernelTimer t(c);
    n0_uside_kernel<<<(n_l1 + 127) / 128, 128, 0, st>>>(P, c->s0.as<double>(), c->s1.as<double>(), c->n0_rowflag.as<int>(),
                                                         c->cl_rows.as<ClRow>(), c->n0_utab.as<URow>());
    n0_kernel<<<dim3(bx, n_chunk), 128, 0, st>>>(P, c->n0_L.as<double>(), c->n0_phi.as<double>(), c->n0_utab.as<URow>(),
                                                 c->cl_rows.as<ClRow>(), c->n0_partial.as<double>());
    n0_reduce_kernel<<<n_L, 128, 0, st>>>(P, c->n0_L.as<double>(), c->n0_partial.as<double>(), c->n0_out.as<double>());
    t.launched(3);
    std::vector<double> host((size_t)6 * n_L);
    LBG_CUDA(cudaMemcpyAsync(host.data(), c->n0_out.p, host.size() * 8, cudaMemcpyDeviceToHost, st));
    int rc = t.finish();
    if (rc) return rc;
    memcpy(out_nkk, host.data(), (size_t)n_L * 8);
    if (out_terms) memcpy(out_terms, host.data() + n_L, (size_t)5 * n_L * 8);
    if (nodes_executed) *nodes_executed = npairs * active_rows;
    return LBG_OK;
}
