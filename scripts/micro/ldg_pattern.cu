// Micro-benchmark: SM-clk per warp load instruction on the L1-hit path for the access patterns of the N0 table gather.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ldg_pattern ldg_pattern.cu
#include <cstdio>
#include <cuda_runtime.h>
// PAT 0: lanes 2k, 2k+1 share a 16-byte element (the N0 gather); 1: one element per lane (coalesced); 2: all lanes one element;
// 3: lanes 4k..4k+3 share; W = bytes per lane (16 or 8)
template <int PAT, int W>
__global__ void k(const char* __restrict__ tab, int iters, double* out) {
    const int lane = threadIdx.x & 31;
    const int e = PAT == 0 ? (lane >> 1) : PAT == 1 ? lane : PAT == 2 ? 0 : (lane >> 2);
    const char* p = tab + (size_t)((blockIdx.x * 37) % 1024) * 256 + (size_t)e * W;
    double ax = 0, ay = 0;
    for (int i = 0; i < iters; ++i) {
        const char* q = p + ((i * 7) & 1023) * 16;
#pragma unroll
        for (int s = 0; s < 8; ++s) {
            if (W == 16) { const double2 v = __ldg(reinterpret_cast<const double2*>(q + s * 262144)); ax += v.x; ay += v.y; }
            else         { const double v = __ldg(reinterpret_cast<const double*>(q + s * 262144)); ax += v; }
        }
    }
    if (ax + ay == 123.456) out[0] = ax;
}
template <int PAT, int W> void run(const char* tab, double* out, const char* name) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int iters = 2000, blocks = 148 * 16, threads = 128;
    float best = 1e9;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0);
        k<PAT, W><<<blocks, threads>>>(tab, iters, out);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep && ms < best) best = ms;
    }
    const double warp_loads = (double)blocks * (threads / 32) * iters * 8;
    printf("%-34s %2d B/lane  %.3f ms  %.2f SM-clk per warp load\n", name, W, best, best * 1e-3 * 1.965e9 * 148 / warp_loads);
}
int main() {
    char* tab; double* out;
    cudaMalloc(&tab, 262144 * 9 + 1024 * 256 + 65536); cudaMemset(tab, 0, 262144 * 9 + 1024 * 256 + 65536); cudaMalloc(&out, 8);
    run<0, 16>(tab, out, "pairs share an element");   run<0, 8>(tab, out, "pairs share an element");
    run<3, 16>(tab, out, "quads share an element");   run<3, 8>(tab, out, "quads share an element");
    run<1, 16>(tab, out, "one element per lane");     run<1, 8>(tab, out, "one element per lane");
    run<2, 16>(tab, out, "all lanes one element");    run<2, 8>(tab, out, "all lanes one element");
    return 0;
}
