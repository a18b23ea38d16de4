// halofit.cu -- HALOFIT spectral parameters (k_nl, n_eff, curvature) for every redshift of a z grid on the device, and the
// P(k, z) rows built from them without a host round trip.  Replaces modi/halofit.py:70-98 (`_get_spec`: two quintic
// InterpolatedUnivariateSpline fits per redshift + an inverse evaluation + two derivatives) as modi/cosmology.py:171-179 applies
// it at every z node.
//
//   sigma_0^2(R_i), 500 nodes            sigma2_kernel (limber.cu), once per linear table
//   R_nl(z): ln sigma^2(R, z) = 0        ln sigma^2(R, z) = ln D(z)^2 + ln sigma_0^2(R), and an interpolating spline commutes with a
//                                        shift of its abscissae, so ONE inverse fit at z = 0 (host, FITPACK order, fitpack.cuh)
//                                        evaluated at -ln D^2 gives ln R_nl(z) (1e-14 against the reference's per-z inverse fits)
//   n_eff, C at ln(1/R_nl)               first / second derivative of the quintic spline through ln(D^2 sigma_0^2(R_i)): fitted PER
//                                        REDSHIFT in FITPACK's own operation order -- the second derivative amplifies the fit's
//                                        rounding by ~1e4, so only the same statements give the same C (fitpack.cuh).  The rotations
//                                        depend on the ln R grid alone: factored once on the host, applied by one device thread per z.
//   P(k, z) rows                         pnl_kernel (limber.cu)
// Redshifts whose R_nl leaves the tabulated decade (the reference's `except` branch) are flagged and left to the caller.
#include "common.cuh"
#include "pk_kernels.cuh"
#include "fitpack.cuh"
#include <cmath>
#include <algorithm>

namespace lbg {
#define HF_MAX_R 512
#define HF_T 32          // redshifts per block: one per lane of warp 0
#define HF_THREADS 256   // the other warps help with staging and the logarithms
// One LANE per redshift, 32 redshifts per block, one block per SM.  The fit is one long dependent chain per redshift (3000 plane
// rotations of the right-hand side, then 500 back-substitution rows with a division each): its time is the latency of the chain, so
// every chain of the call has to be in flight at once.  The control flow of fitpack::fit_apply depends on the abscissae only -- the
// same for every redshift -- so 32 redshifts run in lock step on the lanes of ONE warp at the price of one (a warp per redshift, eight
// per block, needed four waves of blocks for the 4000 redshifts of a forecast: 360 us; this layout needs one).  The factorisation
// (rotations 48 KB, band 24 KB, knots, intervals, sigma_0^2) is staged in shared memory once per block and read by broadcast; the
// right-hand sides live there node-major, y[i][lane] (conflict-free), 4 KB per redshift.  All threads take the logarithms
// y_i = ln(D^2 sigma_0^2(R_i)); warp 0 then runs fit_apply's statements in their order: the six rows a data point rotates into, and
// the five coefficients a back-substitution row reads, live in a register window that slides with the knot interval (the rows leaving
// it are final); of the coefficients only the six the derivative needs are kept.
__global__ void __launch_bounds__(HF_THREADS)
halofit_spec_kernel(int n_rows, int m, const double* __restrict__ s0, const double* __restrict__ tk_g, const int* __restrict__ lint_g,
                    const double* __restrict__ rot_g, const double* __restrict__ band_g, const double* __restrict__ inv_t, const double* __restrict__ inv_c,
                    double lo, double hi, double lnr0, double lnr1, const double* __restrict__ in /*[n][4]: D2, omz, ovz, use*/,
                    HaloZ* __restrict__ rows, int* __restrict__ n_flagged) {
    using namespace fitpack;
    extern __shared__ double hsm[];
    double* rot = hsm;                               // [m][K1][2]
    double* band = rot + 12 * (size_t)m;             // [m][K1]
    double* tk = band + 6 * (size_t)m;               // [m + K1]
    double* s0s = tk + (m + K1);                     // [m]
    double* rinv = s0s + m;                          // [m]: RN(1 / band[i][0]) for the divisions of the back-substitution
    double* yz = rinv + m;                           // [m][HF_T]: y, then z in place
    int* lint = reinterpret_cast<int*>(yz + (size_t)HF_T * m);   // [m]
    __shared__ double sD2[HF_T];
    __shared__ int suse[HF_T];
    __shared__ int sreg[2];
    for (int i = threadIdx.x; i < 12 * m; i += blockDim.x) { const double v = rot_g[i]; rot[i] = (!(i & 1) && v > 1.5) ? 1.0 : v; }   // skipped rotation -> identity
    for (int i = threadIdx.x; i < 6 * m; i += blockDim.x) band[i] = band_g[i];
    for (int i = threadIdx.x; i < m; i += blockDim.x) rinv[i] = __drcp_rn(band_g[(size_t)i * K1]);
    for (int i = threadIdx.x; i < m + K1; i += blockDim.x) tk[i] = tk_g[i];
    for (int i = threadIdx.x; i < m; i += blockDim.x) { s0s[i] = s0[i]; lint[i] = lint_g[i]; }
    const int lane = threadIdx.x;                    // meaningful for warp 0 only
    const int r = blockIdx.x * HF_T + lane;
    const bool mine = threadIdx.x < HF_T && r < n_rows;
    HaloZ h;
    double rknl = 1.0, x = 0.0;
    if (threadIdx.x < HF_T) { sD2[lane] = 1.0; suse[lane] = 0; }
    if (mine) {
        const double* q = in + (size_t)r * 4;
        h.D2 = q[0]; h.rknl = 1.0; h.neff = 0.0; h.ncur = 0.0; h.omz = q[1]; h.ovz = q[2]; h.use_halofit = q[3]; h.pad = 0.0;
        if (q[3] != 0.0) {
            // R_nl from the z = 0 inverse fit at -ln D^2 (modi/halofit.py:71-72 with the abscissae shifted by ln D^2)
            const double tt = -log(q[0]);
            const double lnR = splev(inv_t, m + K1, inv_c, tt);
            rknl = 1.0 / exp(lnR);
            x = log(1.0 / rknl);
            if (!(tt > lo && tt < hi && lnR > lnr0 && lnR < lnr1)) {       // outside the tabulated decade: the caller's exception path
                h.use_halofit = 0.0; h.pad = 1.0; h.rknl = rknl;
                atomicAdd(n_flagged, 1);
            }
        }
        sD2[lane] = q[0];
        suse[lane] = h.use_halofit != 0.0;
    }
    __syncthreads();
    if (threadIdx.x == HF_T) {                       // [sreg[0], sreg[1]): the data points that move the row window by exactly one
        int lo = 1;
        while (lo < m && lint[lo] != lint[lo - 1] + 1) ++lo;
        int hi = lo;
        while (hi < m && lint[hi] == lint[hi - 1] + 1) ++hi;
        sreg[0] = lo; sreg[1] = hi;
    }
    for (int e = threadIdx.x; e < m * HF_T; e += blockDim.x) {              // y[i][t] = ln(D2_t sigma_0^2(R_i)), every thread of the block
        const int t = e & (HF_T - 1), i = e / HF_T;
        if (suse[t]) yz[e] = log(sD2[t] * s0s[i]);
    }
    __syncthreads();
    if (threadIdx.x >= HF_T) return;
    if (mine && h.use_halofit != 0.0) {
        double* y = yz + lane;                       // element i at y[i * HF_T]
      {
        const int nk1 = m;
        double* z = y;                               // z[base] is written after y[base] was consumed (base <= it); same stride
        // ---- right-hand side through the rotations (fpcurf) ----
        // Data point p rotates its right-hand side into the six rows base_p .. base_p + 5, one after the other.  Between points sreg[0]
        // and sreg[1] (interior of the knot vector) base_p = base_(p-1) + 1, and rotation (p, i) needs only (p, i - 1) [the right-hand
        // side] and (p - 1, i + 1) [the row]: rotations 0,1 of point p, 2,3 of point p - 1 and 4,5 of point p - 2 are independent of
        // each other.  They are issued together -- a three-stage software pipeline over the points -- so the dependent chain per point is
        // two rotations deep instead of six; every operation is the one fpcurf does, only the interleaving differs.  Four rows are live
        // at a time (q0..q3; the row that leaves is final and goes to z), operands of the next step are fetched before the chains of
        // this one.  Rotations fpcurf skips (zero pivot) were staged as (cos, sin) = (1, 0), which leaves both operands as they are.
        double w0 = 0, w1 = 0, w2 = 0, w3 = 0, w4 = 0, w5 = 0;       // z[base .. base + 5] of the point-by-point form
        int base = lint[0] - K1;
        const double2* rp = reinterpret_cast<const double2*>(rot);
#define LBG_ROT2(W, Y, CS) { const double s1 = Y, s2 = W; W = add(mul(CS.x, s2), mul(CS.y, s1)); Y = sub(mul(CS.x, s1), mul(CS.y, s2)); }
#define LBG_GENERIC(IT) { const int l = lint[IT]; double yi = y[(size_t)(IT) * HF_T]; const double2* g2 = rp + (size_t)(IT) * K1; \
                          const double2 r0 = g2[0], r1 = g2[1], r2 = g2[2], r3 = g2[3], r4 = g2[4], r5 = g2[5]; \
                          while (base < l - K1) { z[(size_t)base * HF_T] = w0; w0 = w1; w1 = w2; w2 = w3; w3 = w4; w4 = w5; w5 = 0.0; ++base; } \
                          LBG_ROT2(w0, yi, r0) LBG_ROT2(w1, yi, r1) LBG_ROT2(w2, yi, r2) LBG_ROT2(w3, yi, r3) LBG_ROT2(w4, yi, r4) LBG_ROT2(w5, yi, r5) }
        // one pipeline step: stage A = rotations 0,1 of point P, stage B = 2,3 of point P - 1, stage C = 4,5 of point P - 2
#define LBG_PIPE(P, DA, DB, DC) { \
            const int pa = ::min((P) + 1, m - 1), pb = ::min((P), m - 1), pc = ::min(::max((P) - 1, 0), m - 1); \
            const double2 ta0 = rp[(size_t)pa * K1], ta1 = rp[(size_t)pa * K1 + 1], tb2 = rp[(size_t)pb * K1 + 2], tb3 = rp[(size_t)pb * K1 + 3]; \
            const double2 tc4 = rp[(size_t)pc * K1 + 4], tc5 = rp[(size_t)pc * K1 + 5]; \
            const double ynx = y[(size_t)pa * HF_T]; \
            q3 = nx3; nx3 = nx4; nx4 = 0.0; \
            if (DA) LBG_ROT2(q0, yA, a0) if (DB) LBG_ROT2(q1, yB, b2) if (DC) LBG_ROT2(q2, yC, c4) \
            if (DA) LBG_ROT2(q1, yA, a1) if (DB) LBG_ROT2(q2, yB, b3) if (DC) LBG_ROT2(q3, yC, c5) \
            z[(size_t)bb * HF_T] = q0; ++bb; q0 = q1; q1 = q2; q2 = q3; \
            yC = yB; yB = yA; yA = ynx; a0 = ta0; a1 = ta1; b2 = tb2; b3 = tb3; c4 = tc4; c5 = tc5; }
        int it = 0;
        const int reg_lo = sreg[0], reg_hi = sreg[1];
        if (reg_hi - reg_lo >= 3) {
            for (; it < reg_lo; ++it) LBG_GENERIC(it)
            // point reg_lo moves the window by one row: row `base` is final; rows base + 1 .. base + 5 enter the pipeline
            z[(size_t)base * HF_T] = w0;
            double q0 = w1, q1 = w2, q2 = w3, q3 = 0.0, nx3 = w4, nx4 = w5;
            int bb = base + 1;
            double yA = y[(size_t)reg_lo * HF_T], yB = 0.0, yC = 0.0;
            double2 a0 = rp[(size_t)reg_lo * K1], a1 = rp[(size_t)reg_lo * K1 + 1], b2 = a0, b3 = a0, c4 = a0, c5 = a0;
            LBG_PIPE(reg_lo, 1, 0, 0)
            LBG_PIPE(reg_lo + 1, 1, 1, 0)
#pragma unroll 2
            for (int p = reg_lo + 2; p < reg_hi; ++p) LBG_PIPE(p, 1, 1, 1)
            LBG_PIPE(reg_hi, 0, 1, 1)
            LBG_PIPE(reg_hi + 1, 0, 0, 1)
            // back to the point-by-point form: the window of point reg_hi - 1 (its first three rows were written out above)
            base = bb - 3;
            w0 = z[(size_t)base * HF_T]; w1 = z[(size_t)(base + 1) * HF_T]; w2 = z[(size_t)(base + 2) * HF_T];
            w3 = q0; w4 = q1; w5 = q2;
            it = reg_hi;
        }
#pragma unroll 1
        for (; it < m; ++it) LBG_GENERIC(it)
#undef LBG_PIPE
#undef LBG_GENERIC
#undef LBG_ROT2
        z[(size_t)base * HF_T] = w0; z[(size_t)(base + 1) * HF_T] = w1; z[(size_t)(base + 2) * HF_T] = w2;
        z[(size_t)(base + 3) * HF_T] = w3; z[(size_t)(base + 4) * HF_T] = w4; z[(size_t)(base + 5) * HF_T] = w5;
        // ---- interval of the derivative point (spalde) ----
        bool ok = !(x < tk[K1 - 1] || x > tk[nk1]);
        int l = K1;
        while (!(x < tk[l]) && l != nk1) ++l;
        ok = ok && tk[l - 1] < tk[l];
        const int lk = l - K1;                                        // coefficients c[lk .. lk + 5] (0-based) are needed
        // ---- back-substitution (fpback), window c[i+1 .. i+5] ----
        // store / a0 is taken as the correctly rounded quotient from the correctly rounded reciprocal y = RN(1 / a0) (Markstein: q0 = RN(store y),
        // two exact-remainder corrections r = fma(-a0, q, store), q <- fma(r, y, q); the second starts from a faithful q and returns
        // RN(store / a0), the result of the reference's division; tests/test_fitpack_cpu.py) -- 5 dependent operations instead of the ~15
        // of the division sequence, which sits on the chain of every row.  Row i - 1 is fetched before row i's chain, as above; c_i
        // overwrites z_i (read for the last time by its own row) and the six coefficients the derivative needs are read back afterwards.
        double c1 = 0, c2 = 0, c3 = 0, c4 = 0, c5 = 0, keep[K1];
        const double* ar = band + (size_t)(nk1 - 1) * K1;
        double a1n = ar[1], a2n = ar[2], a3n = ar[3], a4n = ar[4], a5n = ar[5], a0n = ar[0], rvn = rinv[nk1 - 1], zn = z[(size_t)(nk1 - 1) * HF_T];
        double a0, a1, a2, a3, a4, a5, rv, store;
#define LBG_ROW(I) { a0 = a0n; a1 = a1n; a2 = a2n; a3 = a3n; a4 = a4n; a5 = a5n; rv = rvn; store = zn; \
                     if ((I) > 0) { const double* an = band + (size_t)((I) - 1) * K1; a0n = an[0]; a1n = an[1]; a2n = an[2]; a3n = an[3]; \
                                    a4n = an[4]; a5n = an[5]; rvn = rinv[(I) - 1]; zn = z[(size_t)((I) - 1) * HF_T]; } }
#define LBG_QUOT(I) { double ci = __dmul_rn(store, rv); ci = __fma_rn(__fma_rn(-a0, ci, store), rv, ci); ci = __fma_rn(__fma_rn(-a0, ci, store), rv, ci); \
                      z[(size_t)(I) * HF_T] = ci; c5 = c4; c4 = c3; c3 = c2; c2 = c1; c1 = ci; }
        int i = nk1 - 1;
        for (; i >= 0 && i > nk1 - 1 - (K1 - 1); --i) {               // the first rows have fewer than five coefficients above them
            LBG_ROW(i)
            const int avail = nk1 - 1 - i;                            // i1 of fpback = min(5, avail)
            if (avail >= 1) store = sub(store, mul(c1, a1));
            if (avail >= 2) store = sub(store, mul(c2, a2));
            if (avail >= 3) store = sub(store, mul(c3, a3));
            if (avail >= 4) store = sub(store, mul(c4, a4));
            LBG_QUOT(i)
        }
#pragma unroll 2
        for (; i >= 0; --i) {
            LBG_ROW(i)
            store = sub(sub(sub(sub(sub(store, mul(c1, a1)), mul(c2, a2)), mul(c3, a3)), mul(c4, a4)), mul(c5, a5));
            LBG_QUOT(i)
        }
#undef LBG_QUOT
#undef LBG_ROW
        for (int e = 0; e < K1; ++e) keep[e] = z[(size_t)(lk + e) * HF_T];
        // ---- derivatives at x (fpader on the six kept coefficients: spalde with the interval found above) ----
        double d[K1] = {0, 0, 0, 0, 0, 0};
        if (ok) {
            double hh[K1 + 1];
            for (int i = 1; i <= K1; ++i) hh[i] = keep[i - 1];
            int kj = K1;
            double fac = 1.0;
            for (int j = 1; j <= K1; ++j) {
                int ki = kj;
                if (j != 1) {
                    int i = K1;
                    for (int jj = j; jj <= K1; ++jj) {
                        const int li = i + lk, lj = li + kj;
                        hh[i] = dvd(sub(hh[i], hh[i - 1]), sub(tk[lj - 1], tk[li - 1]));
                        --i;
                    }
                }
                double dd[K1 + 1];
                for (int i = j; i <= K1; ++i) dd[i] = hh[i];
                if (j != K1) {
                    for (int jj = j + 1; jj <= K1; ++jj) {
                        --ki;
                        int i = K1;
                        for (int j2 = jj; j2 <= K1; ++j2) {
                            const int li = i + lk, lj = li + ki;
                            dd[i] = dvd(add(mul(sub(x, tk[li - 1]), dd[i]), mul(sub(tk[lj - 1], x), dd[i - 1])), sub(tk[lj - 1], tk[li - 1]));
                            --i;
                        }
                    }
                }
                d[j - 1] = mul(dd[K1], fac);
                fac = mul(fac, (double)(K1 - j));
                --kj;
            }
        }
        h.rknl = rknl;
        h.neff = ok ? -d[1] - 3.0 : nan("");
        h.ncur = ok ? -d[2] : nan("");
      }
    }
    if (mine) rows[r] = h;
}
}  // namespace lbg

using namespace lbg;

namespace {
struct HfGrid {                     // per process: the ln R grid the reference fits on, factored once (host) 
    std::mutex mu;
    std::vector<double> lnr;
    fitpack::FitFactor F;
};
HfGrid g_hf;
}

struct lbg_hf_state {               // per context: device copies of the factor, sigma_0^2 on the grid for the current linear table
    std::vector<double> lnr, R, lns0_rev, lnr_rev, inv;
    unsigned long long pk_gen = ~0ull;      // linear table the inverse fit belongs to
    unsigned long long s0_gen = ~0ull;      // linear table whose sigma_0^2 sums were enqueued last
    lbg::DevBuf dev;                // [s0 m | t m+6 | rot 12 m | band 6 m | lint m (int, padded) | inverse fit: t m+6, c m]
    lbg::DevBuf Rdev, zin, rows;    // R grid; build inputs (+ flag counter); HaloZ rows
    lbg::PinBuf pin, pin_in;        // sigma_0^2 read-back (+ flag count); build inputs (own buffers: the build may still be reading
                                    // them when the next call on the context fills the shared staging blocks)
    cudaEvent_t ev_s0a = nullptr, ev_s0 = nullptr, ev_b0 = nullptr, ev_b1 = nullptr;
    bool dev_ok = false, pending = false;
    int pending_launches = 0;
    int init() {
        for (cudaEvent_t* e : {&ev_s0a, &ev_s0, &ev_b0, &ev_b1}) LBG_CUDA(cudaEventCreate(e));
        LBG_CUDA(cudaEventRecord(ev_s0a, 0)); LBG_CUDA(cudaEventRecord(ev_s0, 0));
        return LBG_OK;
    }
    void release() {
        dev.release(); Rdev.release(); zin.release(); rows.release(); pin.release(); pin_in.release();
        for (cudaEvent_t e : {ev_s0a, ev_s0, ev_b0, ev_b1}) if (e) cudaEventDestroy(e);
    }
};

// Waits for the pending build, adds its device time (sigma^2 interval + build interval) and launches to the context's totals.
static int hf_wait_locked(lbg_ctx* c, lbg_hf_state& H, int* n_flagged) {
    cudaError_t e = cudaStreamSynchronize(c->stream);
    if (e == cudaSuccess) e = cudaGetLastError();
    H.pending = false;
    if (e != cudaSuccess) { lbg::set_error("kernel execution failed: %s", cudaGetErrorString(e)); return LBG_ECUDA; }
    float ms0 = 0.f, ms1 = 0.f;
    cudaEventElapsedTime(&ms0, H.ev_s0a, H.ev_s0);
    cudaEventElapsedTime(&ms1, H.ev_b0, H.ev_b1);
    if (!(ms0 >= 0.f)) ms0 = 0.f;
    c->last_ms = (double)ms0 + ms1;
    c->last_launches = H.pending_launches;
    c->total_ms += c->last_ms;
    c->total_launches += H.pending_launches;
    H.pending_launches = 0;
    cudaEventRecord(H.ev_s0a, c->stream);          // a later wait without new sums adds nothing
    cudaEventRecord(H.ev_s0, c->stream);
    *n_flagged = *reinterpret_cast<int*>(H.pin.as<double>() + H.lnr.size());
    return LBG_OK;
}

void lbg_hf_release(lbg_ctx* c) { if (c->hf) { c->hf->release(); delete c->hf; c->hf = nullptr; } }

// modi/halofit.py:70-98 on the host, FITPACK order: lnr increasing, lnsig = ln sigma^2 on it.  out = (rknl = k_nl = 1 / R_nl, neff, ncur).
// *status: 0 ok; 1 = the derivative point ln(1 / rknl) lies outside the fitted range (scipy raises there: the reference's except branch).
// rknl_in > 0: the inverse fit is skipped and that R_nl is used (the refit of the except branch, :82-96).
extern "C" int lbg_halofit_spec_host(int n, const double* lnr, const double* lnsig, double rknl_in, double* out, int* status) {
    LBG_REQUIRE(lnr && lnsig && out && status && n >= 7 && n <= 100000, "lbg_halofit_spec_host: bad argument");
    fitpack::FitFactor inv, fwd;
    std::vector<double> c(n);
    double rknl = rknl_in;
    if (!(rknl_in > 0.0)) {
        std::vector<double> xr(n), yr(n);
        for (int i = 0; i < n; ++i) { xr[i] = lnsig[n - 1 - i]; yr[i] = lnr[n - 1 - i]; }
        LBG_REQUIRE(inv.build(n, xr.data()), "lbg_halofit_spec_host: ln sigma^2 must be strictly decreasing in R (scipy: x must be strictly increasing)");
        inv.solve(yr.data(), c.data());
        rknl = 1.0 / exp(fitpack::splev(inv.t.data(), inv.n, c.data(), 0.0));
    }
    LBG_REQUIRE(fwd.build(n, lnr), "lbg_halofit_spec_host: ln R must be strictly increasing");
    fwd.solve(lnsig, c.data());
    double d[6];
    out[0] = rknl;
    if (!fitpack::spalde(fwd.t.data(), fwd.n, c.data(), log(1.0 / rknl), d)) { *status = 1; out[1] = out[2] = nan(""); return LBG_OK; }
    out[1] = -d[1] - 3.0;
    out[2] = -d[2];
    *status = 0;
    return LBG_OK;
}

// ---- device-side build of the P(k,z) tables -------------------------------------------------------------------------------------
// Pieces (all under the context lock, on the context's stream):
//   hf_ensure_grid     factorisation of the ln R grid on the device (once per grid)
//   hf_enqueue_sigma2  sigma_0^2(R_i) of the current linear table -> pinned host memory, NOT waited for (lbg_pk_prepare_hf: a caller
//                      issues it right after the table upload and does host work meanwhile)
//   hf_finish_sigma2   waits for it, fits the z = 0 inverse spline on the host (50 us; a single device thread would take 10x longer:
//                      3000 dependent Givens rotations with a division and a square root each), uploads it
//   then one spectral-parameter launch for all rows of all tables and one row launch per table.
static int hf_ensure_grid(lbg_ctx* c, lbg_hf_state& H, int m, const double* lnr, const char* who) {
    cudaStream_t st = c->stream;
    const bool grid_same = (int)H.lnr.size() == m && !memcmp(H.lnr.data(), lnr, (size_t)m * 8);
    if (grid_same && H.dev_ok) return LBG_OK;
    std::lock_guard<std::mutex> gg(g_hf.mu);
    if ((int)g_hf.lnr.size() != m || memcmp(g_hf.lnr.data(), lnr, (size_t)m * 8)) {
        LBG_REQUIRE(g_hf.F.build(m, lnr), "%s: ln R grid must be strictly increasing", who);
        g_hf.lnr.assign(lnr, lnr + m);
    }
    const size_t nd = (size_t)m + (m + 6) + 12 * (size_t)m + 6 * (size_t)m;    // then lint (m ints, padded to m doubles), then the inverse fit (2 m + 6)
    if (H.dev.reserve((nd + m + 2 * (size_t)m + 6) * 8 + 64) || H.pin.reserve(((size_t)m + 8) * 8)) return LBG_ENOMEM;
    double* d = H.dev.as<double>();
    LBG_CUDA(cudaMemcpyAsync(d + m, g_hf.F.t.data(), (size_t)(m + 6) * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + m + (m + 6), g_hf.F.rot.data(), (size_t)12 * m * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + m + (m + 6) + 12 * (size_t)m, g_hf.F.a.data(), (size_t)6 * m * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + nd, g_hf.F.lint.data(), (size_t)m * 4, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaStreamSynchronize(st));
    H.lnr.assign(lnr, lnr + m);
    H.dev_ok = true;
    H.pk_gen = ~0ull;
    H.s0_gen = ~0ull;
    return LBG_OK;
}

static int hf_enqueue_sigma2(lbg_ctx* c, lbg_hf_state& H, int m, const double* R, int* launches) {
    if (H.s0_gen == c->pk_gen) return LBG_OK;                 // already in flight / done for this linear table
    cudaStream_t st = c->stream;
    if (H.Rdev.reserve((size_t)m * 8)) return LBG_ENOMEM;
    if (H.R.size() != (size_t)m || memcmp(H.R.data(), R, (size_t)m * 8)) {
        H.R.assign(R, R + m);
        LBG_CUDA(cudaMemcpyAsync(H.Rdev.p, H.R.data(), (size_t)m * 8, cudaMemcpyHostToDevice, st));
    }
    double* d_s0 = H.dev.as<double>();
    const double* kd = c->pk_lnk.as<double>();
    const size_t smem = ((size_t)c->n_k + c->pk_npan0 + c->pk_npan1) * 8;
    if (c->pk_npan0 > 0 && smem <= 45 * 1024 && c->pk_npan0 <= 2048 && c->pk_npan1 <= 2048) {   // the reference's own summation order (pk_kernels.cuh); <= 2 x 31 tree nodes
        const double* sp = c->pk_simps.as<double>();
        sigma2_np_kernel<<<m, 256, smem, st>>>(c->n_k, kd, sp, reinterpret_cast<const SimpsPanel*>(sp + c->n_k), c->pk_npan0, c->pk_npan1,
                                               c->pk_first_dx, c->pk_last_dx, H.Rdev.as<double>(), d_s0);
    } else {
        sigma2_kernel<<<m, 256, 0, st>>>(c->n_k, kd, kd + 2 * (size_t)c->n_k, H.Rdev.as<double>(), d_s0);
    }
    ++*launches;
    LBG_CUDA(cudaMemcpyAsync(H.pin.p, d_s0, (size_t)m * 8, cudaMemcpyDeviceToHost, st));
    LBG_CUDA(cudaEventRecord(H.ev_s0, st));
    H.s0_gen = c->pk_gen;
    return LBG_OK;
}

static int hf_finish_sigma2(lbg_ctx* c, lbg_hf_state& H, int m, const double* lnr, const char* who) {
    if (H.pk_gen == c->pk_gen) return LBG_OK;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaEventSynchronize(H.ev_s0));
    const double* s0 = H.pin.as<double>();
    H.lns0_rev.resize(m); H.lnr_rev.resize(m);
    bool mono = true;
    for (int i = 0; i < m; ++i) { H.lns0_rev[i] = log(s0[m - 1 - i]); H.lnr_rev[i] = lnr[m - 1 - i]; }
    for (int i = 1; i < m; ++i) mono = mono && H.lns0_rev[i] > H.lns0_rev[i - 1];
    LBG_REQUIRE(mono, "%s: sigma^2(R) of this table is not strictly decreasing on the grid", who);
    fitpack::FitFactor inv;
    inv.build(m, H.lns0_rev.data());
    H.inv.resize(2 * (size_t)m + 6);
    memcpy(H.inv.data(), inv.t.data(), (size_t)(m + 6) * 8);
    inv.solve(H.lnr_rev.data(), H.inv.data() + (m + 6));
    const size_t nd = (size_t)m + (m + 6) + 12 * (size_t)m + 6 * (size_t)m;
    LBG_CUDA(cudaMemcpyAsync(H.dev.as<double>() + nd + m, H.inv.data(), H.inv.size() * 8, cudaMemcpyHostToDevice, st));   // pageable: staged before return
    H.pk_gen = c->pk_gen;
    return LBG_OK;
}

// sigma_0^2 on the HALOFIT grid for the linear table just installed, enqueued and not waited for: lbg_pk_build_tables_hf picks the
// result up.  Optional (the build issues it itself when it was not): it lets the caller overlap its own host work with the sums.
extern "C" int lbg_pk_prepare_hf(lbg_ctx* c, int n_R, const double* lnr, const double* R) {
    LBG_REQUIRE(c && lnr && R, "lbg_pk_prepare_hf: NULL argument");
    LBG_REQUIRE(n_R >= 7 && n_R <= HF_MAX_R, "lbg_pk_prepare_hf: n_R=%d outside [7, %d]", n_R, HF_MAX_R);
    std::lock_guard<std::mutex> g(c->mu);
    if (c->n_k <= 0) { set_error("lbg_pk_prepare_hf: call lbg_upload_pk_linear first"); return LBG_ESTATE; }
    LBG_CUDA(cudaSetDevice(c->device));
    if (!c->hf) { c->hf = new lbg_hf_state(); if (int rc = c->hf->init()) return rc; }
    lbg_hf_state& H = *c->hf;
    if (int rc = hf_ensure_grid(c, H, n_R, lnr, "lbg_pk_prepare_hf")) return rc;
    int launches = 0;
    if (H.s0_gen != c->pk_gen) LBG_CUDA(cudaEventRecord(H.ev_s0a, c->stream));
    if (int rc = hf_enqueue_sigma2(c, H, n_R, R, &launches)) return rc;
    H.pending_launches += launches;
    return LBG_OK;
}

// Device-side build of the P(k,z) tables of several z grids in one call.  Table i has n_rows[i] rows and goes to table slot
// slots[i] (lbg_pk_select_slot numbering; -1 = the current slot); zin = the tables' rows one after another,
// zin[row][4] = {D(z)^2, Omega_m(z), Omega_L(z), use_halofit (0 -> linear row)};
// lnr / R: the reference's fitting grid (np.linspace(ln 0.1, ln 10, 500) and its exp, modi/halofit.py:36-37) as the caller's numpy
// computed them.  out_spec (may be NULL): [total rows][4] = {k_nl = 1 / R_nl (the reference's `rknl`), n_eff, curvature, flag}; flag 1 =
// R_nl outside the tabulated decade (the caller evaluates that row by the reference's exception path and passes it through
// lbg_pk_build_table).  *n_flagged receives the number of flagged rows.
// build: 0 = spectral parameters only; 1 = build the tables when no row is flagged, wait for them; 2 = enqueue the table build and
// RETURN WITHOUT WAITING (*n_flagged = -1): the caller does host work meanwhile, any later call on this context is ordered behind
// the build, and lbg_pk_build_wait() reports the flagged count afterwards (> 0: the tables hold linear rows where HALOFIT rows were
// flagged and must be rebuilt through the explicit route before their results are used).  out_spec / out_pnl need build != 2.
// out_pnl (may be NULL, n_tables == 1 only): the rows of the table.
extern "C" int lbg_pk_build_tables_hf(lbg_ctx* c, int n_tables, const int* n_rows_each, const int* slots, const double* zin,
                                      int n_R, const double* lnr, const double* R, double fnu, double om0, double w, int takahashi,
                                      int build, double* out_spec, int* n_flagged, double* out_pnl) {
    LBG_REQUIRE(c && n_rows_each && slots && zin && lnr && R && n_flagged && n_tables > 0 && n_tables <= lbg_ctx::PK_SLOTS,
                "lbg_pk_build_tables_hf: bad argument");
    LBG_REQUIRE(n_R >= 7 && n_R <= HF_MAX_R, "lbg_pk_build_tables_hf: n_R=%d outside [7, %d]", n_R, HF_MAX_R);
    LBG_REQUIRE(out_pnl == nullptr || n_tables == 1, "lbg_pk_build_tables_hf: out_pnl needs n_tables == 1");
    LBG_REQUIRE(build >= 0 && build <= 2 && (build != 2 || (!out_spec && !out_pnl)), "lbg_pk_build_tables_hf: build must be 0, 1 or 2 (2: no host outputs)");
    int n_rows = 0;
    for (int i = 0; i < n_tables; ++i) {
        LBG_REQUIRE(n_rows_each[i] > 0 && slots[i] >= -1 && slots[i] < lbg_ctx::PK_SLOTS, "lbg_pk_build_tables_hf: table %d: bad size / slot", i);
        for (int j = 0; j < i; ++j) LBG_REQUIRE(slots[i] != slots[j], "lbg_pk_build_tables_hf: slot %d given twice", slots[i]);
        n_rows += n_rows_each[i];
    }
    std::lock_guard<std::mutex> g(c->mu);
    if (c->n_k <= 0) { set_error("lbg_pk_build_tables_hf: call lbg_upload_pk_linear first"); return LBG_ESTATE; }
    LBG_CUDA(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    if (!c->hf) { c->hf = new lbg_hf_state(); if (int rc = c->hf->init()) return rc; }
    lbg_hf_state& H = *c->hf;
    if (H.pending) { set_error("lbg_pk_build_tables_hf: an asynchronous build is pending: call lbg_pk_build_wait first"); return LBG_ESTATE; }
    const int m = n_R;
    int launches = 0;
    if (int rc = hf_ensure_grid(c, H, m, lnr, "lbg_pk_build_tables_hf")) return rc;
    if (H.s0_gen != c->pk_gen) LBG_CUDA(cudaEventRecord(H.ev_s0a, st));
    if (int rc = hf_enqueue_sigma2(c, H, m, R, &launches)) return rc;
    if (int rc = hf_finish_sigma2(c, H, m, lnr, "lbg_pk_build_tables_hf")) return rc;
    const size_t nd = (size_t)m + (m + 6) + 12 * (size_t)m + 6 * (size_t)m;
    double* d = H.dev.as<double>();
    const double* d_s0 = d;
    const double* d_t = d + m;
    const double* d_rot = d_t + (m + 6);
    const double* d_band = d_rot + 12 * (size_t)m;
    const int* d_lint = reinterpret_cast<const int*>(d + nd);
    const double* d_inv = d + nd + m;
    // ---- per row on the device: R_nl from the inverse fit, flags, the per-redshift forward fit ----
    const size_t in_bytes = (size_t)n_rows * 32;
    if (H.pin_in.reserve(in_bytes + 64) || H.zin.reserve(in_bytes + 64) || H.rows.reserve((size_t)n_rows * sizeof(HaloZ)) ||
        (out_spec && c->pin_out.reserve((size_t)n_rows * sizeof(HaloZ) + 64)))
        return LBG_ENOMEM;
    memcpy(H.pin_in.p, zin, in_bytes);
    LBG_CUDA(cudaEventRecord(H.ev_b0, st));
    LBG_CUDA(cudaMemcpyAsync(H.zin.p, H.pin_in.p, in_bytes, cudaMemcpyHostToDevice, st));
    int* d_flag = reinterpret_cast<int*>(H.zin.as<unsigned char>() + in_bytes);
    LBG_CUDA(cudaMemsetAsync(d_flag, 0, 4, st));
    const size_t hsmem = (12 * (size_t)m + 6 * (size_t)m + (m + 6) + 2 * (size_t)m + (size_t)HF_T * m) * 8 + (size_t)m * 4;
    static bool attr_set = false;                  // the attribute is per function, not per device context
    if (!attr_set) {
        LBG_CUDA(cudaFuncSetAttribute(halofit_spec_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      (int)(((21 + HF_T) * (size_t)HF_MAX_R + 6) * 8 + HF_MAX_R * 4)));
        attr_set = true;
    }
    halofit_spec_kernel<<<(n_rows + HF_T - 1) / HF_T, HF_THREADS, hsmem, st>>>(n_rows, m, d_s0, d_t, d_lint, d_rot, d_band, d_inv, d_inv + (m + 6),
                                                           H.lns0_rev[0], H.lns0_rev[m - 1], lnr[0], lnr[m - 1], H.zin.as<double>(),
                                                           H.rows.as<HaloZ>(), d_flag);
    ++launches;
    int* hflag = reinterpret_cast<int*>(H.pin.as<double>() + m);
    LBG_CUDA(cudaMemcpyAsync(hflag, d_flag, 4, cudaMemcpyDeviceToHost, st));
    int flagged = 0;
    if (build == 1) {                              // the decision to build needs the count
        LBG_CUDA(cudaStreamSynchronize(st));
        flagged = *hflag;
    }
    const bool do_build = build == 2 || (build == 1 && flagged == 0);
    const int n_k = c->n_k;
    if (do_build) {
        HaloConst hc{fnu, om0, w, takahashi};
        const double* kd = c->pk_lnk.as<double>();
        int row0 = 0;
        for (int i = 0; i < n_tables; ++i) {
            const int nr = n_rows_each[i];
            const bool cur = slots[i] < 0 || slots[i] == c->pk_cur;
            DevBuf& buf = cur ? c->pk_pnl : c->pk_store[slots[i]];
            if (buf.reserve((size_t)nr * n_k * 8)) return LBG_ENOMEM;
            pnl_kernel<<<dim3((n_k + 127) / 128, nr), 128, 0, st>>>(nr, n_k, kd, kd + n_k, H.rows.as<HaloZ>() + row0, hc, buf.as<double>());
            ++launches;
            if (out_pnl) LBG_CUDA(cudaMemcpyAsync(out_pnl, buf.p, (size_t)nr * n_k * 8, cudaMemcpyDeviceToHost, st));
            row0 += nr;
        }
        for (int i = 0; i < n_tables; ++i) {
            if (slots[i] < 0 || slots[i] == c->pk_cur) c->pnl_nz = n_rows_each[i]; else c->store_nz[slots[i]] = n_rows_each[i];
        }
    }
    if (out_spec) LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, H.rows.p, (size_t)n_rows * sizeof(HaloZ), cudaMemcpyDeviceToHost, st));
    LBG_CUDA(cudaEventRecord(H.ev_b1, st));
    LBG_CUDA(cudaGetLastError());
    H.pending_launches += launches;
    H.pending = true;
    if (build == 2) { *n_flagged = -1; return LBG_OK; }
    int rc = hf_wait_locked(c, H, &flagged);
    if (rc) return rc;
    if (out_spec) {
        const HaloZ* hz = c->pin_out.as<HaloZ>();
        for (int r = 0; r < n_rows; ++r) {
            out_spec[4 * r] = hz[r].rknl; out_spec[4 * r + 1] = hz[r].neff; out_spec[4 * r + 2] = hz[r].ncur; out_spec[4 * r + 3] = hz[r].pad;
        }
    }
    *n_flagged = flagged;
    return LBG_OK;
}

// Waits for the build lbg_pk_build_tables_hf(build = 2) enqueued; *n_flagged = rows whose R_nl left the tabulated decade (0: the tables
// are good).  Without a pending build: *n_flagged = 0.
extern "C" int lbg_pk_build_wait(lbg_ctx* c, int* n_flagged) {
    LBG_REQUIRE(c && n_flagged, "lbg_pk_build_wait: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    *n_flagged = 0;
    if (!c->hf || !c->hf->pending) return LBG_OK;
    LBG_CUDA(cudaSetDevice(c->device));
    return hf_wait_locked(c, *c->hf, n_flagged);
}

// Number of k nodes of the linear table installed on the context (0: none): callers size their host buffers for
// lbg_pk_build_table / lbg_pk_build_table_hf output with it.
extern "C" int lbg_pk_table_nk(lbg_ctx* c, int* n_k) {
    LBG_REQUIRE(c && n_k, "lbg_pk_table_nk: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    *n_k = c->n_k;
    return LBG_OK;
}
