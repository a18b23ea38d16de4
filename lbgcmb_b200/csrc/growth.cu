// growth.cu -- un-normalised linear growth D+(a) tables for many cosmologies at once.
//
// Replaces the 1000 scipy `romberg` calls of modi/cosmology.py:92-111 (`_eval_Dplus`):
//     y(a_i) = E(a_i) * romberg(kernel, ln a_0, ln a_i),   kernel(ln a) = (a E(a))^-3 a,
//     E(a) = a^-1.5 sqrt(M + (1 - M - L) a + L a^3)                                   (modi/cosmology.py:119-125)
// One thread per (cosmology, node); every thread runs the scipy-1.2.1 Romberg rule as the reference's host code does
// (scipy/integrate/quadrature.py of 1.2.1: trapezoid refinements by _difftrap, Richardson table, stop when the last two diagonal
// entries differ by less than tol or rtol |result|, at most 2^10 panels), including numpy's pairwise order for the sum of the new
// ordinates of a refinement, so a node stops at the same level as on the host and agrees to rounding (1e-14, tests/test_growth_gpu.py).
// A cosmology-parameter sweep needs one table per cosmology (0.11 s each in the reference's host code); here 64 cosmologies x
// 1000 nodes are one launch.
#include "common.cuh"
#include <cmath>

namespace lbg {

__device__ __forceinline__ double growth_E(double a, double M, double L) {
    return pow(a, -1.5) * sqrt(M + (1.0 - M - L) * a + L * pow(a, 3.0));
}
__device__ __forceinline__ double growth_kernel(double loga, double M, double L) {
    const double a = exp(loga);
    return pow(a * growth_E(a, M, L), -3.0) * a;
}

// numpy's pairwise sum of f(lox + h j), j = 0..n-1, for n a power of two <= 512 (numpy/core/src/umath/loops_utils.h pairwise_sum:
// n < 8 sequential from 0; n <= 128 eight running sums combined as ((r0+r1)+(r2+r3))+((r4+r5)+(r6+r7)); larger n split in halves).
__device__ double block_sum_128(double lox, double h, int j0, int n, double M, double L) {
    if (n < 8) {
        double res = 0.0;
        for (int j = 0; j < n; ++j) res += growth_kernel(lox + h * (double)(j0 + j), M, L);
        return res;
    }
    double r[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) r[k] = growth_kernel(lox + h * (double)(j0 + k), M, L);
    for (int i = 8; i < n; i += 8) {
#pragma unroll
        for (int k = 0; k < 8; ++k) r[k] += growth_kernel(lox + h * (double)(j0 + i + k), M, L);
    }
    return ((r[0] + r[1]) + (r[2] + r[3])) + ((r[4] + r[5]) + (r[6] + r[7]));
}
__device__ double pairwise_sum(double lox, double h, int n, double M, double L) {
    if (n <= 128) return block_sum_128(lox, h, 0, n, M, L);
    if (n == 256) return block_sum_128(lox, h, 0, 128, M, L) + block_sum_128(lox, h, 128, 128, M, L);
    return (block_sum_128(lox, h, 0, 128, M, L) + block_sum_128(lox, h, 128, 128, M, L)) +
           (block_sum_128(lox, h, 256, 128, M, L) + block_sum_128(lox, h, 384, 128, M, L));          // n == 512
}

__global__ void growth_table_kernel(int n_cosmo, int n, const double* __restrict__ Ms, const double* __restrict__ Ls,
                                    const double* __restrict__ loga, double tol, double rtol, double* __restrict__ out,
                                    int* __restrict__ levels) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x, ic = blockIdx.y;
    if (i >= n || ic >= n_cosmo) return;
    const double M = Ms[ic], L = Ls[ic];
    const double a = loga[0], b = loga[i];
    const double span = b - a;
    double acc = 0.5 * (growth_kernel(a, M, L) + growth_kernel(b, M, L));
    double prev[11], cur[11];
    prev[0] = span * acc;
    double result = prev[0];
    int panels = 1, level = 1;
    for (; level <= 10; ++level) {
        panels *= 2;
        const int half = panels / 2;
        const double h = span / (double)half;
        acc = acc + pairwise_sum(a + 0.5 * h, h, half, M, L);
        cur[0] = span * acc / (double)panels;
        double p4 = 1.0;
        for (int k = 0; k < level; ++k) {
            p4 *= 4.0;
            cur[k + 1] = (p4 * cur[k] - prev[k]) / (p4 - 1.0);
        }
        result = cur[level];
        const double err = fabs(result - prev[level - 1]);
        if (err < tol || err < rtol * fabs(result)) break;
        for (int k = 0; k <= level; ++k) prev[k] = cur[k];
    }
    out[(size_t)ic * n + i] = growth_E(exp(b), M, L) * result;
    if (levels) levels[(size_t)ic * n + i] = level;
}
}  // namespace lbg

using namespace lbg;

extern "C" int lbg_growth_table(lbg_ctx* c, int n_cosmo, const double* omega_m, const double* omega_l, int n, const double* loga,
                                double* out_y /*[n_cosmo][n]*/, int* out_levels /*[n_cosmo][n] or NULL*/) {
    LBG_REQUIRE(c && omega_m && omega_l && loga && out_y, "lbg_growth_table: NULL argument");
    LBG_REQUIRE(n_cosmo > 0 && n_cosmo <= 65535 && n > 0, "lbg_growth_table: bad sizes n_cosmo=%d n=%d", n_cosmo, n);
    for (int i = 1; i < n; ++i) LBG_REQUIRE(loga[i] > loga[i - 1], "lbg_growth_table: ln a nodes must be increasing");
    for (int i = 0; i < n_cosmo; ++i) LBG_REQUIRE(omega_m[i] > 0, "lbg_growth_table: omega_m[%d] must be positive", i);
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t nc = (size_t)n_cosmo * 8, nn = (size_t)n * 8, no = (size_t)n_cosmo * n * 8;
    if (c->s0.reserve(2 * nc) || c->s1.reserve(nn) || c->s2.reserve(no) || c->s3.reserve(no / 2)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    double* dm = c->s0.as<double>();
    LBG_CUDA(cudaMemcpyAsync(dm, omega_m, nc, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(dm + n_cosmo, omega_l, nc, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, loga, nn, cudaMemcpyHostToDevice, st));
    KernelTimer t(c);
    growth_table_kernel<<<dim3((n + 63) / 64, n_cosmo), 64, 0, st>>>(n_cosmo, n, dm, dm + n_cosmo, c->s1.as<double>(), 1.48e-8, 1.48e-8,
                                                                    c->s2.as<double>(), out_levels ? c->s3.as<int>() : nullptr);
    t.launched();
    LBG_CUDA(cudaStreamSynchronize(st));
    LBG_CUDA(cudaMemcpyAsync(out_y, c->s2.p, no, cudaMemcpyDeviceToHost, st));
    if (out_levels) LBG_CUDA(cudaMemcpyAsync(out_levels, c->s3.p, no / 2, cudaMemcpyDeviceToHost, st));
    return t.finish();
}
