// xchg.cu -- the exchange step of a sharded forecast over NVLink peer memory: every rank's block of doubles
// [partial sums | its shard of the curves] lands in every other rank's memory with plain peer stores, one kernel per rank, no
// intermediate library call.  The sums are then added on the host in rank order, which makes the reduced values independent of
// the collective algorithm (bit-reproducible from run to run and from N to N for a given sharding).
//
// Layout of one rank's window (one cudaMalloc, exported with cudaIpcGetMemHandle and opened by the peers):
//   flags[world]            uint64, flags[q] = number of the last exchange whose block from rank q has arrived here
//   err                     int, set by the kernel when a peer did not arrive within the time limit
//   staging[cap]            this rank's outgoing block (H2D target)
//   slots[2][world][cap]    incoming blocks, double-buffered by the parity of the exchange number
// Exchange k: CTA p of rank r copies staging -> slots[k & 1][r] of PEER p (16-byte peer stores over NVLink), fences, raises
// flags[r] on peer p (release, system scope), then waits for flags[p] >= k in its OWN window (acquire).  No rank reads or writes a
// peer's memory while waiting, so a late peer costs nothing but the wait.  Rank r can only begin exchange k + 2 -- the next one to
// reuse the same slots -- after every peer has started exchange k + 1, i.e. after that peer's device-to-host copy of exchange k was
// issued before its kernel k + 1 on the same stream: two buffers are enough.
// A peer that never arrives (crashed process) ends the wait after ~4 s of GPU clocks with an error instead of hanging the GPU.
#include "common.cuh"
#include <cstring>

#define LBG_XCHG_MAX_WORLD 16

struct lbg_xchg_state {
    int rank = 0, world = 0;
    size_t cap = 0;                              // doubles per block
    unsigned char* base = nullptr;               // own window
    unsigned char* peer[LBG_XCHG_MAX_WORLD] = {nullptr};
    bool opened[LBG_XCHG_MAX_WORLD] = {false};
    unsigned long long epoch = 0;
    lbg::PinBuf pin_in, pin_out;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    float last_ms = 0.f;
};

namespace lbg {
struct XchgParams {
    int rank, world;
    unsigned long long epoch, limit;
    size_t cap, n;                               // block capacity and payload, in doubles
    size_t off_stage, off_slots;                 // byte offsets inside a window
    unsigned char* win[LBG_XCHG_MAX_WORLD];      // win[rank] = own window
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long* p) {
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}

__global__ void __launch_bounds__(256) xchg_kernel(XchgParams P) {
    const int p = blockIdx.x;                                            // the peer this CTA serves
    const double* src = reinterpret_cast<const double*>(P.win[P.rank] + P.off_stage);
    double* dst = reinterpret_cast<double*>(P.win[p] + P.off_slots) + ((P.epoch & 1ull) * P.world + P.rank) * P.cap;
    const size_t n2 = P.n >> 1;
    const double2* s2 = reinterpret_cast<const double2*>(src);
    double2* d2 = reinterpret_cast<double2*>(dst);                       // cap is even and the regions are 256-byte aligned
    for (size_t i = threadIdx.x; i < n2; i += blockDim.x) d2[i] = s2[i];
    if ((P.n & 1) && threadIdx.x == 0) dst[P.n - 1] = src[P.n - 1];
    __syncthreads();                                                     // every store of the block issued ...
    if (threadIdx.x == 0) {
        __threadfence_system();                                          // ... and ordered before the flag, system-wide
        st_release_sys(reinterpret_cast<unsigned long long*>(P.win[p]) + P.rank, P.epoch);
        const unsigned long long* mine = reinterpret_cast<const unsigned long long*>(P.win[P.rank]) + p;
        const long long t0 = clock64();
        while (ld_acquire_sys(mine) < P.epoch) {
            if ((unsigned long long)(clock64() - t0) > P.limit) {
                *reinterpret_cast<int*>(P.win[P.rank] + LBG_XCHG_MAX_WORLD * 8) = 1 + p;
                break;
            }
        }
    }
}
}  // namespace lbg

using namespace lbg;

static const size_t XCHG_OFF_ERR = LBG_XCHG_MAX_WORLD * 8, XCHG_OFF_STAGE = 256;

void lbg_xchg_release(lbg_ctx* c) {
    lbg_xchg_state* X = c->xchg;
    if (!X) return;
    for (int q = 0; q < X->world; ++q)
        if (X->opened[q]) cudaIpcCloseMemHandle(X->peer[q]);
    if (X->base) cudaFree(X->base);
    X->pin_in.release(); X->pin_out.release();
    if (X->e0) cudaEventDestroy(X->e0);
    if (X->e1) cudaEventDestroy(X->e1);
    delete X;
    c->xchg = nullptr;
}

extern "C" int lbg_xchg_create(lbg_ctx* c, int rank, int world, long long cap_doubles, unsigned char* out_handle /*[64]*/) {
    LBG_REQUIRE(c && out_handle, "lbg_xchg_create: NULL argument");
    LBG_REQUIRE(world >= 2 && world <= LBG_XCHG_MAX_WORLD && rank >= 0 && rank < world, "lbg_xchg_create: rank %d of %d (2 .. %d ranks)", rank, world, LBG_XCHG_MAX_WORLD);
    LBG_REQUIRE(cap_doubles >= 2 && cap_doubles <= (1LL << 24), "lbg_xchg_create: block capacity %lld outside [2, 2^24] doubles", cap_doubles);
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    lbg_xchg_release(c);
    lbg_xchg_state* X = new lbg_xchg_state;
    c->xchg = X;
    X->rank = rank; X->world = world;
    X->cap = ((size_t)cap_doubles + 31) & ~(size_t)31;                   // blocks stay 256-byte aligned
    const size_t bytes = XCHG_OFF_STAGE + (1 + 2 * (size_t)world) * X->cap * 8;
    LBG_CUDA(cudaMalloc(reinterpret_cast<void**>(&X->base), bytes));
    LBG_CUDA(cudaMemset(X->base, 0, bytes));
    LBG_CUDA(cudaDeviceSynchronize());
    LBG_CUDA(cudaEventCreate(&X->e0));
    LBG_CUDA(cudaEventCreate(&X->e1));
    cudaIpcMemHandle_t h;
    LBG_CUDA(cudaIpcGetMemHandle(&h, X->base));
    memcpy(out_handle, &h, 64);
    X->peer[rank] = X->base;
    return LBG_OK;
}

extern "C" int lbg_xchg_open(lbg_ctx* c, const unsigned char* handles /*[world][64], rank order*/) {
    LBG_REQUIRE(c && handles, "lbg_xchg_open: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    lbg_xchg_state* X = c->xchg;
    if (!X) { set_error("lbg_xchg_open: call lbg_xchg_create first"); return LBG_ESTATE; }
    LBG_CUDA(cudaSetDevice(c->device));
    for (int q = 0; q < X->world; ++q) {
        if (q == X->rank || X->opened[q]) continue;
        cudaIpcMemHandle_t h;
        memcpy(&h, handles + (size_t)q * 64, 64);
        void* p = nullptr;
        LBG_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        X->peer[q] = reinterpret_cast<unsigned char*>(p);
        X->opened[q] = true;
    }
    return LBG_OK;
}

// One exchange of the n doubles already packed in X->pin_in: on return X->pin_out holds [world][n] in rank order.  Caller holds c->mu.
static int xchg_run(lbg_ctx* c, lbg_xchg_state* X, size_t n) {
    const size_t nb = n * 8;
    cudaStream_t st = c->stream;
    ++X->epoch;
    XchgParams P{};
    P.rank = X->rank; P.world = X->world; P.epoch = X->epoch; P.limit = 8000000000ull;
    P.cap = X->cap; P.n = n; P.off_stage = XCHG_OFF_STAGE; P.off_slots = XCHG_OFF_STAGE + X->cap * 8;
    for (int q = 0; q < X->world; ++q) P.win[q] = X->peer[q];
    LBG_CUDA(cudaMemcpyAsync(X->base + XCHG_OFF_STAGE, X->pin_in.p, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaEventRecord(X->e0, st));
    xchg_kernel<<<X->world, 256, 0, st>>>(P);
    LBG_CUDA(cudaEventRecord(X->e1, st));
    const unsigned char* slots = X->base + P.off_slots + (X->epoch & 1ull) * X->world * X->cap * 8;
    LBG_CUDA(cudaMemcpy2DAsync(X->pin_out.p, nb, slots, X->cap * 8, nb, X->world, cudaMemcpyDeviceToHost, st));
    int* herr = reinterpret_cast<int*>(X->pin_out.as<unsigned char>() + nb * X->world);
    LBG_CUDA(cudaMemcpyAsync(herr, X->base + XCHG_OFF_ERR, 4, cudaMemcpyDeviceToHost, st));
    LBG_CUDA(cudaStreamSynchronize(st));
    if (*herr) {
        set_error("lbg_xchg: rank %d did not arrive within the time limit (exchange %llu)", *herr - 1, X->epoch);
        cudaMemset(X->base + XCHG_OFF_ERR, 0, 4);
        return LBG_ECUDA;
    }
    LBG_CUDA(cudaGetLastError());
    LBG_CUDA(cudaEventElapsedTime(&X->last_ms, X->e0, X->e1));
    return LBG_OK;
}

static int xchg_ready(lbg_ctx* c, lbg_xchg_state** out, size_t n) {
    lbg_xchg_state* X = c->xchg;
    if (!X) { set_error("lbg_xchg: call lbg_xchg_create / lbg_xchg_open first"); return LBG_ESTATE; }
    for (int q = 0; q < X->world; ++q)
        if (!X->peer[q]) { set_error("lbg_xchg: window of rank %d not opened", q); return LBG_ESTATE; }
    LBG_REQUIRE(n >= 1 && n <= X->cap, "lbg_xchg: block of %zu doubles outside [1, %zu]", n, X->cap);
    LBG_CUDA(cudaSetDevice(c->device));
    if (X->pin_in.reserve(n * 8) || X->pin_out.reserve(n * 8 * X->world + 64)) return LBG_ENOMEM;
    *out = X;
    return LBG_OK;
}

// All ranks call this the same number of times with the same n.  block[n] (host) -> out[world][n] (host), rank order.
extern "C" int lbg_xchg_allgather(lbg_ctx* c, long long n, const double* block, double* out, double* out_ms /*or NULL*/) {
    LBG_REQUIRE(c && block && out && n >= 1, "lbg_xchg_allgather: NULL argument or empty block");
    std::lock_guard<std::mutex> g(c->mu);
    lbg_xchg_state* X = nullptr;
    if (int rc = xchg_ready(c, &X, (size_t)n)) return rc;
    memcpy(X->pin_in.p, block, (size_t)n * 8);
    if (int rc = xchg_run(c, X, (size_t)n)) return rc;
    if (out_ms) *out_ms = X->last_ms;
    memcpy(out, X->pin_out.p, (size_t)n * 8 * X->world);
    return LBG_OK;
}

// The exchange step of a sharded forecast in one call: sums[n_sums] are added over the ranks (in rank order) and the curves --
// curves[n_curves][n_mine], this rank's cyclic shard (units rank, rank + world, ...) of n units -- come back whole, out_curves[n_curves][n].
extern "C" int lbg_xchg_sums_curves(lbg_ctx* c, int n_sums, const double* sums, int n_curves, long long n, const double* curves,
                                    double* out_sums, double* out_curves, double* out_ms /*or NULL*/) {
    LBG_REQUIRE(c && n_sums >= 0 && n_curves >= 0 && n >= 0 && n_sums + n_curves > 0, "lbg_xchg_sums_curves: bad sizes");
    LBG_REQUIRE((n_sums == 0 || (sums && out_sums)) && (n_curves == 0 || n == 0 || out_curves), "lbg_xchg_sums_curves: NULL argument");
    std::lock_guard<std::mutex> g(c->mu);
    if (!c->xchg) { set_error("lbg_xchg: call lbg_xchg_create / lbg_xchg_open first"); return LBG_ESTATE; }
    const int world = c->xchg->world, rank = c->xchg->rank;
    const size_t width = ((size_t)n + world - 1) / world;
    const size_t n_mine = (size_t)n > (size_t)rank ? ((size_t)n - rank + world - 1) / world : 0;
    LBG_REQUIRE(n_mine == 0 || curves || n_curves == 0, "lbg_xchg_sums_curves: NULL curves");
    const size_t nd = (size_t)n_sums + (size_t)n_curves * width;
    lbg_xchg_state* X = nullptr;
    if (int rc = xchg_ready(c, &X, nd > 0 ? nd : 1)) return rc;
    double* in = X->pin_in.as<double>();
    if (nd == 0) in[0] = 0.0;
    for (int i = 0; i < n_sums; ++i) in[i] = sums[i];
    for (int q = 0; q < n_curves; ++q) {
        double* row = in + n_sums + (size_t)q * width;
        if (n_mine) memcpy(row, curves + (size_t)q * n_mine, n_mine * 8);
        for (size_t j = n_mine; j < width; ++j) row[j] = 0.0;
    }
    const size_t ne = nd > 0 ? nd : 1;
    if (int rc = xchg_run(c, X, ne)) return rc;
    if (out_ms) *out_ms = X->last_ms;
    const double* got = X->pin_out.as<double>();
    for (int i = 0; i < n_sums; ++i) {
        double acc = got[i];
        for (int r = 1; r < world; ++r) acc += got[(size_t)r * ne + i];          // rank order
        out_sums[i] = acc;
    }
    for (int r = 0; r < world; ++r) {
        const size_t m = (size_t)n > (size_t)r ? ((size_t)n - r + world - 1) / world : 0;
        for (int q = 0; q < n_curves; ++q) {
            const double* src = got + (size_t)r * ne + n_sums + (size_t)q * width;
            double* dst = out_curves + (size_t)q * n + r;
            for (size_t j = 0; j < m; ++j) dst[j * world] = src[j];
        }
    }
    return LBG_OK;
}

extern "C" int lbg_xchg_destroy(lbg_ctx* c) {
    LBG_REQUIRE(c, "lbg_xchg_destroy: ctx is NULL");
    std::lock_guard<std::mutex> g(c->mu);
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    lbg_xchg_release(c);
    return LBG_OK;
}
