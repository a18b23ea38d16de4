// fisher_dev.cuh -- device helpers shared by fisher.cu and sweep.cu: fixed-order block sums, the per-multipole Gaussian
// covariance entry (interloper.py:42-63) and the in-thread equilibrated LU solve.
#pragma once
#include <cmath>

namespace lbg {

constexpr int RED_T = 256;

// Block-wide sum of `nv` values per thread (nv <= 32 per pass); result valid in thread 0.
template <int NV>
__device__ __forceinline__ void block_sum(double (&v)[NV], double* sm /*[NV][RED_T]*/) {
#pragma unroll
    for (int q = 0; q < NV; ++q) sm[q * RED_T + threadIdx.x] = v[q];
    __syncthreads();
    for (int h = RED_T / 2; h > 0; h >>= 1) {
        if (threadIdx.x < h) {
#pragma unroll
            for (int q = 0; q < NV; ++q) sm[q * RED_T + threadIdx.x] += sm[q * RED_T + threadIdx.x + h];
        }
        __syncthreads();
    }
#pragma unroll
    for (int q = 0; q < NV; ++q) v[q] = sm[q * RED_T];
    __syncthreads();
}

// (T^{ac} T^{bd} + T^{ad} T^{bc}) / (2 l + 1) / fsky with the reference's operation order (interloper.py:50-58).
__device__ __forceinline__ double gaussian_cov_entry(const double* T, int n_f, int a, int b, int c, int d, double ell, double fsky) {
    double r = __dmul_rn(T[a * n_f + c], T[b * n_f + d]);
    r = __dadd_rn(r, __dmul_rn(T[a * n_f + d], T[b * n_f + c]));
    r = __ddiv_rn(r, __dadd_rn(__dmul_rn(2.0, ell), 1.0));
    return __ddiv_rn(r, fsky);
}

// In-thread solve of the symmetric system C X = B by LU with partial pivoting after symmetric diagonal equilibration
// (C_ij / sqrt(C_ii C_jj)): band-power covariances span many decades between kk and gg, the correlation matrix does not.
template <int MS, int MR>
__device__ void solve_spd(int ns, int nr, double* A /*[ns][ns], destroyed*/, double* B /*[nr][ns] in, solution out*/) {
    double sc[MS];
    for (int i = 0; i < ns; ++i) sc[i] = 1.0 / sqrt(fabs(A[i * ns + i]));
    for (int i = 0; i < ns; ++i)
        for (int j = 0; j < ns; ++j) A[i * ns + j] *= sc[i] * sc[j];
    for (int r = 0; r < nr; ++r)
        for (int i = 0; i < ns; ++i) B[r * ns + i] *= sc[i];
    for (int k = 0; k < ns; ++k) {
        int p = k;
        double best = fabs(A[k * ns + k]);
        for (int i = k + 1; i < ns; ++i) {
            const double v = fabs(A[i * ns + k]);
            if (v > best) { best = v; p = i; }
        }
        if (p != k) {
            for (int j = 0; j < ns; ++j) { const double t = A[k * ns + j]; A[k * ns + j] = A[p * ns + j]; A[p * ns + j] = t; }
            for (int r = 0; r < nr; ++r) { const double t = B[r * ns + k]; B[r * ns + k] = B[r * ns + p]; B[r * ns + p] = t; }
        }
        const double piv = 1.0 / A[k * ns + k];
        for (int i = k + 1; i < ns; ++i) {
            const double f = A[i * ns + k] * piv;
            if (f != 0.0) {
                for (int j = k + 1; j < ns; ++j) A[i * ns + j] -= f * A[k * ns + j];
                for (int r = 0; r < nr; ++r) B[r * ns + i] -= f * B[r * ns + k];
            }
        }
    }
    for (int r = 0; r < nr; ++r)
        for (int i = ns - 1; i >= 0; --i) {
            double s = B[r * ns + i];
            for (int j = i + 1; j < ns; ++j) s -= A[i * ns + j] * B[r * ns + j];
            B[r * ns + i] = s / A[i * ns + i];
        }
    for (int r = 0; r < nr; ++r)
        for (int i = 0; i < ns; ++i) B[r * ns + i] *= sc[i];
}

}  // namespace lbg
