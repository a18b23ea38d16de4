// common.cuh -- context object, error plumbing and small device helpers shared by all kernels.
#pragma once
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>
#include <mutex>
#include "../../include/lbgcmb.h"

#define LBG_PAIR_STRIDE 16384      /* compile-time spectrum stride of cl_pairs: the 32 gathers of a node share one base register */

namespace lbg {

void set_error(const char* fmt, ...);

#define LBG_CUDA(call)                                                                     \
    do {                                                                                   \
        cudaError_t e__ = (call);                                                          \
        if (e__ != cudaSuccess) {                                                          \
            lbg::set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e__)); \
            return LBG_ECUDA;                                                              \
        }                                                                                  \
    } while (0)

#define LBG_REQUIRE(cond, ...)            \
    do {                                  \
        if (!(cond)) {                    \
            lbg::set_error(__VA_ARGS__);  \
            return LBG_EINVAL;            \
        }                                 \
    } while (0)

// Grow-only device buffer.
struct DevBuf {
    void*  p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return LBG_OK;
        if (p) cudaFree(p);
        p = nullptr; cap = 0;
        cudaError_t e = cudaMalloc(&p, bytes);
        if (e != cudaSuccess) { set_error("cudaMalloc(%zu) -> %s", bytes, cudaGetErrorString(e)); return LBG_ENOMEM; }
        cap = bytes;
        return LBG_OK;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
    void release() { if (p) cudaFree(p); p = nullptr; cap = 0; }
};

// Grow-only pinned host staging buffer (true asynchronous DMA; pageable copies are staged by the driver piece by piece).
struct PinBuf {
    void*  p = nullptr;
    size_t cap = 0;
    int reserve(size_t bytes) {
        if (bytes <= cap) return LBG_OK;
        if (p) cudaFreeHost(p);
        p = nullptr; cap = 0;
        bytes = (bytes + (1u << 16)) & ~(size_t)((1u << 16) - 1);
        cudaError_t e = cudaHostAlloc(&p, bytes, cudaHostAllocDefault);
        if (e != cudaSuccess) { set_error("cudaHostAlloc(%zu) -> %s", bytes, cudaGetErrorString(e)); return LBG_ENOMEM; }
        cap = bytes;
        return LBG_OK;
    }
    template <class T> T* as() const { return reinterpret_cast<T*>(p); }
    void release() { if (p) cudaFreeHost(p); p = nullptr; cap = 0; }
};

// One row of the CMB spline table per integer knot: uniform cubic B-spline coefficients of the eight
// spectra, 64 bytes, so that one knot = two 32-byte sectors.
struct __align__(64) ClRow {
    double uTT, uTE, uEE, uBB;   // unlensed (gradient) spectra, enter f_alpha
    double lTT, lTE, lEE, lBB;   // lensed spectra, enter the filters (noise added analytically)
};

}  // namespace lbg

struct lbg_n0_plan;               // n0.cu: grid-dependent part of an N0 call, kept between calls
struct lbg_hf_state;              // halofit.cu: HALOFIT fitting grid / sigma_0^2 state
struct lbg_xchg_state;            // xchg.cu: peer-memory windows of the exchange step

struct lbg_ctx {
    int          device = 0;
    int          lane = 0;
    int          sm_count = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t  ev0 = nullptr, ev1 = nullptr;
    cudaEvent_t  ev2 = nullptr, ev3 = nullptr;   // around the dominant kernel of the last call (roofline timing)
    double       last_main_ms = 0.0;
    double       last_ms = 0.0;
    int          last_launches = 0;
    double       total_ms = 0.0;     // since the last lbg_kernel_totals(reset)
    long long    total_launches = 0;
    std::mutex   mu;
    // CMB tables
    int          n_ell = 0;          // knots 0..n_ell-1
    lbg::DevBuf  cl_rows;            // ClRow[n_ell + 2], row r holds coefficient index r-1
    lbg::DevBuf  cl_soa;             // the same coefficients spectrum-major: double[8][cl_soa_stride]
    int          cl_soa_stride = 0;
    lbg::DevBuf  cl_poly;            // per-call: double[8 slots][4][LBG_PAIR_STRIDE] monomial coefficients per knot interval, noise folded in
    bool         ub_zero = false;    // unlensed BB identically zero (its slot is then never read)
    std::vector<double> cl_host;     // the uploaded tables, [8][n_ell] in ClRow order: value comparison on re-upload, scipy-order splines
    unsigned long long  cl_gen = 0;  // bumped by every upload that changed the tables
    std::vector<double> cl_bspl_t, cl_bspl_c;   // scipy make_interp_spline(k=3) representation (knots; [8][n_ell] coefficients), built on demand
    unsigned long long  cl_bspl_gen = ~0ull;
    // N0 workspaces and state
    lbg::DevBuf  n0_L, n0_utab, n0_partial, n0_out, n0_plan_dev;
    lbg::PinBuf  pin_n0, pin_n0_out;
    lbg_n0_plan* n0_plan = nullptr;
    bool         n0_pending = false; // lbg_n0_ho02_begin issued, lbg_n0_ho02_end not yet
    int          n0_n_L = 0, n0_n_lev = 0;
    int          n0_path = 0;        // integrand kernel of the last call: 5 = range-walking (rows5), 3 = generic per-node masks (rows3), 6 = multi-level (rows5m)
    long long    n0_nodes_walked = 0;
    double       n0_exact_below = 2.0;   // l1 rows below this get their u-side splines from the scipy-order host routine
    std::vector<double> n0_exact_vals;
    unsigned long long  n0_exact_gen = ~0ull;
    bool         cw_lane_valid = false;
    // generic scratch
    lbg::DevBuf  s0, s1, s2, s3, s4, s5;
    lbg::PinBuf  pin_in, pin_out;    // pinned staging for the short, copy-dominated calls (lbg_limber)
    // P(k) / Limber state (see limber.cu)
    int          n_k = 0;
    unsigned long long pk_gen = 0;   // bumped whenever a new linear table / k grid is installed
    lbg_hf_state* hf = nullptr;
    lbg_xchg_state* xchg = nullptr;
    lbg::DevBuf  pk_lnk, pk_plin, pk_rdk;
    lbg::DevBuf  pk_simps;           // numpy-order Simpson data of the ln k grid: [Delta0^2 n_k | SimpsPanel x (npan0 + npan1)]
    int          pk_npan0 = 0, pk_npan1 = 0;
    double       pk_first_dx = 0.0, pk_last_dx = 0.0;
    int          pnl_nz = 0;         // rows of the CURRENT P(k,z) table (pk_pnl); 0 = none built
    lbg::DevBuf  pk_pnl;
    // parked P(k,z) tables (lbg_pk_select_slot): a forecast alternates between a few z grids (C_gg, C_kg, C_kk nodes) and
    // would otherwise rebuild the same rows for every spectrum; all are dropped when a new linear table / k grid is installed
    static const int PK_SLOTS = 4;
    int          pk_cur = 0;
    int          store_nz[PK_SLOTS] = {0, 0, 0, 0};
    lbg::DevBuf  pk_store[PK_SLOTS];
    void drop_pk_tables() { pnl_nz = 0; for (int i = 0; i < PK_SLOTS; ++i) store_nz[i] = 0; ++pk_gen; }
    // spectra kept on the device between calls (lbg_limber_keep -> lbg_field_fisher_sets): set s is a [set_rows][set_n_ell] matrix
    static const int SPEC_SETS = 20;
    lbg::DevBuf  spec_set[SPEC_SETS];
    int          set_rows[SPEC_SETS] = {0}, set_n_ell[SPEC_SETS] = {0};
};

namespace lbg {
int upload_rdk(lbg_ctx* c, int n_k, const double* k);   // limber.cu: reciprocal k-cell widths for the installed k grid
int ensure_scipy_bspline(lbg_ctx* c);                   // ctx.cu: fills cl_bspl_t / cl_bspl_c for the current tables
// scipy.interpolate.make_interp_spline(arange(n), y, k=3) bit for bit (not-a-knot knots, LAPACK dgbsv order): ctx.cu
void scipy_notaknot_cubic(int n, const double* y, double* t_out /*[n+4] or NULL*/, double* c_out /*[n]*/);
struct KernelTimer {   // records ev0/ev1 on the context stream around the kernels of one API call
    lbg_ctx* c;
    explicit KernelTimer(lbg_ctx* ctx) : c(ctx) { c->last_launches = 0; cudaEventRecord(c->ev0, c->stream); }
    void launched(int n = 1) { c->last_launches += n; }
    int finish() {
        cudaError_t e = cudaEventRecord(c->ev1, c->stream);
        if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
        if (e == cudaSuccess) { float ms = 0.f; e = cudaEventElapsedTime(&ms, c->ev0, c->ev1); c->last_ms = ms; c->total_ms += ms; c->total_launches += c->last_launches; }
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) { set_error("kernel execution failed: %s", cudaGetErrorString(e)); return LBG_ECUDA; }
        return LBG_OK;
    }
};
}  // namespace lbg
