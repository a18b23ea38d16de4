// Micro-benchmark: L1 data-stage cost of LDG.E.128 when only a fraction of the lanes is predicated on.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldg_pred ldg_pred.cu && ./ldg_pred
#include <cstdio>
#include <cuda_runtime.h>
template <int MOD>
__global__ void k(const double2* __restrict__ tab, int iters, double* out) {
    const int lane = threadIdx.x & 31;
    const bool on = (lane % MOD) == 0;
    const double2* p = tab + (blockIdx.x * 37 % 1024) * 16 + (lane >> 1);   // neighbouring lanes share an interval, like the N0 gather
    double ax = 0, ay = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int s = 0; s < 8; ++s) {
            if (on) { const double2 v = __ldg(p + s * 16384 + ((i * 7) & 1023)); ax += v.x; ay += v.y; }
        }
    }
    if (ax + ay == 123.456) out[0] = ax;
}
int main() {
    double2* tab; double* out;
    cudaMalloc(&tab, sizeof(double2) * 16384 * 9); cudaMemset(tab, 0, sizeof(double2) * 16384 * 9); cudaMalloc(&out, 8);
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000, blocks = 148 * 16, threads = 128;
    for (int mod = 1; mod <= 4; ++mod) {
        float best = 1e9;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            if (mod == 1) k<1><<<blocks, threads>>>(tab, iters, out);
            if (mod == 2) k<2><<<blocks, threads>>>(tab, iters, out);
            if (mod == 3) k<3><<<blocks, threads>>>(tab, iters, out);
            if (mod == 4) k<4><<<blocks, threads>>>(tab, iters, out);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
        }
        const double warp_loads = (double)blocks * (threads / 32) * iters * 8;
        printf("lanes on: every %d   %.3f ms   %.2f SM-clk per warp-LDG.128\n", mod, best, best * 1e-3 * 1.965e9 * 148 / warp_loads);
    }
    return 0;
}
