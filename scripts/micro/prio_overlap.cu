// Can a short kernel on a high-priority stream get SM slots while a long low-priority grid (register-file filling, 4 CTAs/SM of
// 128 threads x 128 registers) still has CTAs pending?  Depends on the shared-memory carve-out the resident grid left the SMs in.
// nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o prio_overlap prio_overlap.cu ; ./prio_overlap
#include <cstdio>
#include <chrono>
#include <thread>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(128, 4) long_kernel(double* out, int spin) {
    __shared__ double s[128];
    double a = threadIdx.x * 1e-3, b = 1.000001;
    double r[48];
#pragma unroll
    for (int i = 0; i < 48; ++i) r[i] = a + i;
    for (int k = 0; k < spin; ++k) {
#pragma unroll
        for (int i = 0; i < 48; ++i) r[i] = fma(r[i], b, a);
    }
    double t = 0;
#pragma unroll
    for (int i = 0; i < 48; ++i) t += r[i];
    s[threadIdx.x] = t;
    __syncthreads();
    if (t == 123.456) out[blockIdx.x] = s[(threadIdx.x + 1) & 127];
}

__global__ void short_kernel(double* out, int n) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < n; i += blockDim.x) sm[i] = i;
    __syncthreads();
    if (threadIdx.x == 0) out[blockIdx.x] = n ? sm[n - 1] : 1.0;
}

static double now() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); }

int main() {
    int lo, hi;
    cudaDeviceGetStreamPriorityRange(&lo, &hi);
    cudaStream_t sl, sh;
    cudaStreamCreateWithPriority(&sl, cudaStreamNonBlocking, lo);
    cudaStreamCreateWithPriority(&sh, cudaStreamNonBlocking, hi);
    double* d; cudaMalloc(&d, 1 << 24);
    printf("priority range least %d greatest %d\n", lo, hi);
    const int carveA[] = {-1, -1, -1, 50, 50, 50, 50};
    const int smemB[]  = {0, 1024, 17408, 0, 17408, 17408, 17408};
    const int carveB[] = {-1, -1, -1, -1, -1, 50, 100};
    for (int v = 0; v < 7; ++v) {
        cudaFuncSetAttribute(long_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carveA[v]);
        cudaFuncSetAttribute(short_kernel, cudaFuncAttributePreferredSharedMemoryCarveout, carveB[v]);
        for (int rep = 0; rep < 2; ++rep) {
            cudaDeviceSynchronize();
            double t0 = now();
            long_kernel<<<400000, 128, 0, sl>>>(d, 1500);
            std::this_thread::sleep_for(std::chrono::milliseconds(3));
            double t1 = now();
            short_kernel<<<160, 128, smemB[v], sh>>>(d, smemB[v] / 8);
            cudaStreamSynchronize(sh);
            double t2 = now();
            cudaStreamSynchronize(sl);
            double t3 = now();
            if (rep) printf("carveA %3d  B smem %5d carveB %3d : short kernel done %.2f ms after its launch; long grid ends at %.2f ms (err %s)\n",
                            carveA[v], smemB[v], carveB[v], t2 - t1, t3 - t0, cudaGetErrorString(cudaGetLastError()));
        }
    }
    return 0;
}
