// div_rn(a, b, RN(1/b)) == RN(a / b) ?  (the five-instruction quotient of csrc/limber.cu).  Random mantissas and exponents, plus the
// structured cases that break naive schemes (b with an all-ones mantissa, a just below / above b, quotients next to 1 and 2).
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o div_markstein div_markstein.cu ; ./div_markstein
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ double div_rn(double a, double b, double rb) {
    double q = a * rb;
    if (!(fabs(q) > 1e-250 && fabs(q) < 1e250)) return __ddiv_rn(a, b);
    double r = fma(-q, b, a);
    q = fma(r, rb, q);
    r = fma(-q, b, a);
    return fma(r, rb, q);
}
__device__ __forceinline__ double div_one_step(double a, double b, double rb) {   // what NOT to use: one residual step only
    double q = a * rb;
    double r = fma(-q, b, a);
    return fma(r, rb, q);
}
__device__ __forceinline__ uint64_t rng(uint64_t& s) { s ^= s << 13; s ^= s >> 7; s ^= s << 17; return s; }
__device__ __forceinline__ double mk(uint64_t mant, int e) { return __longlong_as_double(((uint64_t)(1023 + e) << 52) | (mant & 0xFFFFFFFFFFFFFull)); }

__global__ void check(unsigned long long* bad2, unsigned long long* bad1, int iters, int mode) {
    uint64_t s = 0x9E3779B97F4A7C15ull * (blockIdx.x * (uint64_t)blockDim.x + threadIdx.x + 1) + mode;
    unsigned long long n2 = 0, n1 = 0;
    for (int i = 0; i < iters; ++i) {
        uint64_t ma = rng(s), mb = rng(s), e = rng(s);
        int ea = (int)(e % 61) - 30, eb = (int)((e >> 8) % 61) - 30;
        if (mode == 1) mb = 0xFFFFFFFFFFFFFull - (mb & 0xFF);          // b mantissa (almost) all ones
        if (mode == 2) { ma = mb + (int64_t)((e >> 20) % 9) - 4; ea = eb; }  // a within a few ulp of b
        if (mode == 3) { mb &= 0xFFFFFull << 32; ma &= 0xFFFFFull << 32; }   // short mantissas: exact and near-exact quotients
        const double a = mk(ma, ea), b = mk(mb, eb);
        const double want = a / b, rb = 1.0 / b;
        n2 += __double_as_longlong(div_rn(a, b, rb)) != __double_as_longlong(want);
        n1 += __double_as_longlong(div_one_step(a, b, rb)) != __double_as_longlong(want);
    }
    atomicAdd(bad2, n2); atomicAdd(bad1, n1);
}

int main() {
    unsigned long long *d, h[2];
    cudaMalloc(&d, 16);
    for (int mode = 0; mode < 4; ++mode) {
        cudaMemset(d, 0, 16);
        check<<<148 * 8, 256>>>(d, d + 1, 3300, mode);
        cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
        printf("mode %d: %.3g pairs, two-step mismatches %llu, one-step mismatches %llu (%s)\n", mode, 148.0 * 8 * 256 * 3300, h[0], h[1],
               cudaGetErrorString(cudaGetLastError()));
    }
    return 0;
}
