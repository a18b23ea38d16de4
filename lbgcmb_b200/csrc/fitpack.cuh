// fitpack.cuh -- the part of Dierckx' FITPACK that scipy's InterpolatedUnivariateSpline(x, y, k=5) runs (fpcurf with s = 0:
// interpolating spline), restated operation by operation so that the coefficients -- and the second derivative HALOFIT takes from
// them -- carry the same rounding as the reference's (modi/halofit.py:70-98).
//
// Why the operation order matters: the curvature C = -d^2 ln sigma^2 / d ln R^2 is the second derivative of a quintic spline through
// 500 samples spaced 0.0092 apart; the ~3000 Givens rotations of the fit leave ~5e-15 of rounding on the coefficients, the second
// derivative multiplies it by ~1/h^2 ~ 1e4, so two correct implementations of the same spline differ by ~1e-10 in C (DESIGN.md
// "parity notes").  The same statements in the same order do not.
//
// Structure used here: the rotations depend on the abscissae only.  `FitFactor` holds, for a fixed x grid, the knots, the sequence of
// (cos, sin) pairs and the triangular band R; `fit_apply` runs the right-hand-side half of fpcurf for one y vector and back-substitutes
// (fpback).  HALOFIT fits ln sigma^2 on the SAME ln R grid at every redshift, so one factorisation serves every z and the per-z work
// (a device thread per redshift, halofit.cu) is 3000 plane rotations of the right-hand side + one back-substitution.
#pragma once
#include <cmath>
#include <vector>

#ifdef __CUDACC__
#define LBG_HD __host__ __device__ __forceinline__
#else
#define LBG_HD inline
#endif

namespace lbg {
namespace fitpack {

// FITPACK is compiled without FMA contraction in the scipy wheels: every product and sum below is rounded on its own.
#ifdef __CUDA_ARCH__
LBG_HD double mul(double a, double b) { return __dmul_rn(a, b); }
LBG_HD double add(double a, double b) { return __dadd_rn(a, b); }
LBG_HD double sub(double a, double b) { return __dsub_rn(a, b); }
LBG_HD double dvd(double a, double b) { return __ddiv_rn(a, b); }
LBG_HD double sqr(double a) { return __dsqrt_rn(a); }
#else
LBG_HD double mul(double a, double b) { volatile double r = a * b; return r; }
LBG_HD double add(double a, double b) { volatile double r = a + b; return r; }
LBG_HD double sub(double a, double b) { volatile double r = a - b; return r; }
LBG_HD double dvd(double a, double b) { volatile double r = a / b; return r; }
LBG_HD double sqr(double a) { return sqrt(a); }
#endif

constexpr int K = 5, K1 = 6;

// fpbspl: the K1 non-zero B-splines of degree K at x, t[l-1] <= x < t[l] in FITPACK's 1-based interval index l
LBG_HD void fpbspl(const double* t, double x, int l, double* h /*[K1]*/) {
    double hh[K1];
    h[0] = 1.0;
    for (int j = 1; j <= K; ++j) {
        for (int i = 0; i < j; ++i) hh[i] = h[i];
        h[0] = 0.0;
        for (int i = 1; i <= j; ++i) {
            const int li = l + i, lj = li - j;              // 1-based
            const double tli = t[li - 1], tlj = t[lj - 1];
            if (tli == tlj) { h[i] = 0.0; continue; }
            const double f = dvd(hh[i - 1], sub(tli, tlj));
            h[i - 1] = add(h[i - 1], mul(f, sub(tli, x)));
            h[i] = mul(f, sub(x, tlj));
        }
    }
}

// splev at one point: value = sum_j c(l-k1+j) h(j).  The interval is the one splev.f's linear walk stops at (the first l >= k1 with
// x < t(l+1), or nk1), found by bisection: same l, 9 steps instead of ~250.
LBG_HD double splev(const double* t, int n, const double* c, double x) {
    const int nk1 = n - K1;
    int l = nk1;
    if (x < t[nk1 - 1]) {
        int lo = K1, hi = nk1 - 1;                          // invariant: the answer lies in [lo, hi], x < t[hi]
        while (lo < hi) { const int mid = (lo + hi) >> 1; if (x < t[mid]) hi = mid; else lo = mid + 1; }
        l = lo;
    }
    double h[K1];
    fpbspl(t, x, l, h);
    double sp = 0.0;
    int ll = l - K1;
    for (int j = 0; j < K1; ++j) { sp = add(sp, mul(c[ll], h[j])); ++ll; }
    return sp;
}

// spalde / fpader: d[j] = j-th derivative at x, j = 0..K.  Returns false when x lies outside [t(k1), t(n-k)] (spalde's ier = 10).
LBG_HD bool spalde(const double* t, int n, const double* c, double x, double* d /*[K1]*/) {
    const int nk1 = n - K1;
    if (x < t[K1 - 1] || x > t[nk1]) return false;
    int l = K1;
    while (!(x < t[l]) && l != nk1) ++l;
    if (t[l - 1] >= t[l]) return false;
    double h[K1 + 1];                                       // 1-based below
    const int lk = l - K1;
    for (int i = 1; i <= K1; ++i) h[i] = c[i + lk - 1];
    int kj = K1;
    double fac = 1.0;
    for (int j = 1; j <= K1; ++j) {
        int ki = kj;
        if (j != 1) {
            int i = K1;
            for (int jj = j; jj <= K1; ++jj) {
                const int li = i + lk, lj = li + kj;
                h[i] = dvd(sub(h[i], h[i - 1]), sub(t[lj - 1], t[li - 1]));
                --i;
            }
        }
        double dd[K1 + 1];
        for (int i = j; i <= K1; ++i) dd[i] = h[i];
        if (j != K1) {
            for (int jj = j + 1; jj <= K1; ++jj) {
                --ki;
                int i = K1;
                for (int j2 = jj; j2 <= K1; ++j2) {
                    const int li = i + lk, lj = li + ki;
                    dd[i] = dvd(add(mul(sub(x, t[li - 1]), dd[i]), mul(sub(t[lj - 1], x), dd[i - 1])), sub(t[lj - 1], t[li - 1]));
                    --i;
                }
            }
        }
        d[j - 1] = mul(dd[K1], fac);
        const double ak = (double)(K1 - j);
        fac = mul(fac, ak);
        --kj;
    }
    return true;
}

// Right-hand-side half of fpcurf for one y vector on a factored grid, then fpback.  rot[it][j] = (cos, sin) or skip (cos = 2) in
// application order; lint[it] = FITPACK's 1-based knot interval of data point it; a = [n_c][K1] triangular band.  z, c: [n_c] scratch / out.
LBG_HD void fit_apply(int m, const int* lint, const double* rot /*[m][K1][2]*/, const double* a /*[m][K1]*/, const double* y,
                      double* z, double* c) {
    const int nc = m;                                       // interpolation: as many coefficients as data points
    for (int i = 0; i < nc; ++i) z[i] = 0.0;
    for (int it = 0; it < m; ++it) {
        double yi = y[it];
        int j = lint[it] - K1;                              // 1-based row index before the increment
        for (int i = 0; i < K1; ++i) {
            ++j;
            const double cs = rot[((size_t)it * K1 + i) * 2], sn = rot[((size_t)it * K1 + i) * 2 + 1];
            if (cs > 1.5) continue;                         // piv == 0: no rotation
            const double stor1 = yi, stor2 = z[j - 1];
            z[j - 1] = add(mul(cs, stor2), mul(sn, stor1));
            yi = sub(mul(cs, stor1), mul(sn, stor2));
        }
    }
    // fpback(a, z, n = nc, k = K1, c)
    const int k1 = K1 - 1;
    c[nc - 1] = dvd(z[nc - 1], a[(size_t)(nc - 1) * K1]);
    int i = nc - 1;                                         // 1-based index of the row being solved
    for (int j = 2; j <= nc; ++j) {
        double store = z[i - 1];
        int i1 = k1;
        if (j <= k1) i1 = j - 1;
        int mm = i;
        for (int l = 1; l <= i1; ++l) {
            ++mm;
            store = sub(store, mul(c[mm - 1], a[(size_t)(i - 1) * K1 + l]));
        }
        c[i - 1] = dvd(store, a[(size_t)(i - 1) * K1]);
        --i;
    }
}

// Left-hand-side half of fpcurf (s = 0, w = 1) for m strictly increasing abscissae: knots, rotations, band.
struct FitFactor {
    int m = 0, n = 0;
    std::vector<double> t, rot, a;
    std::vector<int> lint;
    bool build(int m_, const double* x) {
        m = m_;
        n = m + K1;
        if (m < K1 + 1) return false;
        for (int i = 1; i < m; ++i) if (!(x[i] > x[i - 1])) return false;
        t.assign(n, 0.0);
        // interior knots of the interpolating spline (odd degree): t(k+2 ..) = x(k/2 + 2 ..)
        const int mk1 = m - K1, k3 = K / 2;
        int i = K1 + 1, j = k3 + 2;
        for (int l = 1; l <= mk1; ++l) { t[i - 1] = x[j - 1]; ++i; ++j; }
        for (int q = 1; q <= K1; ++q) { t[q - 1] = x[0]; t[n - K1 + q - 1] = x[m - 1]; }
        const int nk1 = n - K1;
        a.assign((size_t)nk1 * K1, 0.0);
        rot.assign((size_t)m * K1 * 2, 0.0);
        lint.assign(m, 0);
        int l = K1;
        for (int it = 0; it < m; ++it) {
            const double xi = x[it];
            while (!(xi < t[l]) && l != nk1) ++l;
            lint[it] = l;
            double h[K1];
            fpbspl(t.data(), xi, l, h);
            int jr = l - K1;
            for (int q = 0; q < K1; ++q) {
                ++jr;
                const double piv = h[q];
                double* r2 = &rot[((size_t)it * K1 + q) * 2];
                if (piv == 0.0) { r2[0] = 2.0; r2[1] = 0.0; continue; }
                // fpgivs(piv, a(jr,1), cos, sin)
                double& ww = a[(size_t)(jr - 1) * K1];
                const double store = fabs(piv);
                double dd;
                if (store >= ww) { const double q2 = dvd(ww, piv); dd = mul(store, sqr(add(1.0, mul(q2, q2)))); }
                else { const double q2 = dvd(piv, ww); dd = mul(ww, sqr(add(1.0, mul(q2, q2)))); }
                const double cs = dvd(ww, dd), sn = dvd(piv, dd);
                ww = dd;
                r2[0] = cs; r2[1] = sn;
                if (q == K1 - 1) break;
                int i2 = 0;
                for (int i1 = q + 1; i1 < K1; ++i1) {        // fprota(cos, sin, h(i1), a(jr, i2))
                    ++i2;
                    const double stor1 = h[i1], stor2 = a[(size_t)(jr - 1) * K1 + i2];
                    a[(size_t)(jr - 1) * K1 + i2] = add(mul(cs, stor2), mul(sn, stor1));
                    h[i1] = sub(mul(cs, stor1), mul(sn, stor2));
                }
            }
        }
        return true;
    }
    void solve(const double* y, double* c_out) const {
        std::vector<double> z(m);
        fit_apply(m, lint.data(), rot.data(), a.data(), y, z.data(), c_out);
    }
};

}  // namespace fitpack
}  // namespace lbg
