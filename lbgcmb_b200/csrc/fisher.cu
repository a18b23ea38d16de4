// fisher.cu -- per-multipole covariance / S-N / Fisher accumulation (the "L4" stage of the forecast path).
//
// Replaces: snr.py:21-25 (snr), interloper.py:42-63 (get_ClsCov), interloper.py:223-253 (per-L inverse covariance,
//           Fisher and bias-vector sums), yu.py:76-108 (per-L tracer correlation matrix, rho_L^2 = r^T rho^-1 r),
//           mqw/py/mqw.py:91-104 and cadence/cadence.py:156-166 (N_p(z_i) Fisher outwith the Schur limit).
// Every sum over multipoles is a fixed-order tree (block partials -> one final block): deterministic, no atomics.
// The work is tiny at reference sizes (<= 4951 multipoles, matrices <= 20 x 20): these kernels are latency bound and
// no roofline fraction is claimed for them.
#include "common.cuh"
#include <cstring>
#include "fisher_dev.cuh"
#include <cmath>

namespace lbg {

// snr.py:21-25: sum over l <= lmax of nmodes (C / sqrt(Var))^2 ; one block, fixed order.
__global__ void __launch_bounds__(RED_T)
snr_kernel(int n, const double* __restrict__ ell, const double* __restrict__ cl, const double* __restrict__ var,
           const double* __restrict__ nmodes, double lmax, double* __restrict__ out) {
    __shared__ double sm[RED_T];
    double acc[1] = {0.0};
    for (int i = threadIdx.x; i < n; i += RED_T) {
        if (ell[i] <= lmax) {
            const double r = cl[i] / sqrt(var[i]);
            acc[0] += nmodes[i] * (r * r);
        }
    }
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0) out[0] = sqrt(acc[0]);
}

// The experiment x delensing x dilution sweep of snr.py:107-133 in one launch: CTA (d, r, e) evaluates, for every multipole,
//   var = (((1 - rho_r^2) C_kk + N_kk^e) (C_gg + N_gg^d) + C_kg^2) / ((2 L + 1) fsky_e)            (Ckg.py:61-87 with rho, :69-71)
// with numpy's roundings (one per operation, as the var_Ckg mirror produces it on the host), then the sum of snr_kernel above in
// the same order: the value equals var_Ckg(...) followed by snr(...) for that design bit for bit.
__global__ void __launch_bounds__(RED_T)
snr_grid_kernel(int n, const double* __restrict__ ell, const double* __restrict__ nmodes, const double* __restrict__ ckk,
                const double* __restrict__ cgg, const double* __restrict__ ckg, const double* __restrict__ nkk /*[n_exp][n]*/,
                const double* __restrict__ fsky /*[n_exp]*/, const double* __restrict__ rho /*[n_rho][n]*/,
                const double* __restrict__ ngg /*[n_dil]*/, double lmax, double* __restrict__ out /*[n_exp][n_rho][n_dil]*/) {
    __shared__ double sm[RED_T];
    const int d = blockIdx.x, r = blockIdx.y, e = blockIdx.z;
    const double* nk = nkk + (size_t)e * n;
    const double* rh = rho + (size_t)r * n;
    const double fs = fsky[e], ng = ngg[d];
    double acc[1] = {0.0};
    for (int i = threadIdx.x; i < n; i += RED_T) {
        const double L = ell[i];
        if (L <= lmax) {
            const double kk = __dmul_rn(__dsub_rn(1.0, __dmul_rn(rh[i], rh[i])), ckk[i]);
            const double num = __dadd_rn(__dmul_rn(__dadd_rn(kk, nk[i]), __dadd_rn(cgg[i], ng)), __dmul_rn(ckg[i], ckg[i]));
            const double den = __dmul_rn(__dadd_rn(__dmul_rn(2.0, L), 1.0), fs);
            const double var = __ddiv_rn(num, den);
            const double q = ckg[i] / sqrt(var);
            acc[0] += nmodes[i] * (q * q);
        }
    }
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0) out[((size_t)e * gridDim.y + r) * gridDim.x + d] = sqrt(acc[0]);
}

// interloper.py:42-63 for any set of fields: tot[l][a][b] = (C + N)^{ab}(l); spectrum i is the field pair (sa[i], sb[i]).
// cov[l][i][j] = (T^{a c} T^{b d} + T^{a d} T^{b c}) / (2 l + 1) / fsky  with (a,b) = pair i, (c,d) = pair j.
__global__ void gaussian_cov_kernel(int n_ell, int n_f, int n_spec, const double* __restrict__ ell, const double* __restrict__ tot,
                                    const int* __restrict__ sa, const int* __restrict__ sb, double fsky, double* __restrict__ cov) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long n = (long long)n_ell * n_spec * n_spec;
    if (t >= n) return;
    const int j = (int)(t % n_spec), i = (int)((t / n_spec) % n_spec), l = (int)(t / ((long long)n_spec * n_spec));
    const double* T = tot + (size_t)l * n_f * n_f;
    const int a = sa[i], b = sb[i], c = sa[j], d = sb[j];
    cov[t] = gaussian_cov_entry(T, n_f, a, b, c, d, ell[l], fsky);
}

constexpr int MAX_PAR = 8;

// One thread per multipole: X = C^-1 [dD_0 .. dD_{np-1}, dDelta], then q[l] = {dD_p . X_q (np x np), dD_p . X_delta (np)}.
// interloper.py:223-253 (np = 2, ns = 3), yu.py:99-106 (np = 1, ns = number of tracers).
template <int MS>
__global__ void __launch_bounds__(64)
quadform_kernel(int n_ell, int ns, int np, const double* __restrict__ cov, const double* __restrict__ dD,
                const double* __restrict__ dDelta, double* __restrict__ q /*[n_ell][np*np + np]*/) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n_ell) return;
    double A[MS * MS], B[(MAX_PAR + 1) * MS], D[MAX_PAR * MS];
    for (int e = 0; e < ns * ns; ++e) A[e] = cov[(size_t)l * ns * ns + e];
    for (int e = 0; e < np * ns; ++e) { D[e] = dD[(size_t)l * np * ns + e]; B[e] = D[e]; }
    const int nr = np + (dDelta ? 1 : 0);
    if (dDelta) for (int i = 0; i < ns; ++i) B[np * ns + i] = dDelta[(size_t)l * ns + i];
    solve_spd<MS, MAX_PAR + 1>(ns, nr, A, B);
    const int nq = np * np + np;
    double* o = q + (size_t)l * nq;
    for (int p = 0; p < np; ++p) {
        for (int r = 0; r < np; ++r) {
            double s = 0.0;
            for (int i = 0; i < ns; ++i) s += D[p * ns + i] * B[r * ns + i];
            o[p * np + r] = s;
        }
        double s = 0.0;
        if (dDelta) for (int i = 0; i < ns; ++i) s += D[p * ns + i] * B[np * ns + i];
        o[np * np + p] = s;
    }
}

// out[e] = sum_l w[l] q[l][e]; one block per entry e, fixed order.
__global__ void __launch_bounds__(RED_T)
weighted_sum_kernel(int n_ell, int nq, const double* __restrict__ w, const double* __restrict__ q, double* __restrict__ out) {
    __shared__ double sm[RED_T];
    const int e = blockIdx.x;
    double acc[1] = {0.0};
    for (int l = threadIdx.x; l < n_ell; l += RED_T) acc[0] += w[l] * q[(size_t)l * nq + e];
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0) out[e] = acc[0];
}

struct MqwBin { double bp, bs, Np, Ns, wp, ws, wps, pad; };   // per redshift shell, mqw/py/mqw.py:24-41

// mqw.py:63-94: A00(l) = sum_i [cgg (bp Np)^2 + wp], sshift(l) = sum_i A0i^2 / Aii, S = A00 / (A00 - sshift).
__global__ void mqw_schur_kernel(int n_ell, int n_bin, size_t cgg_stride, const double* __restrict__ cgg,
                                 const MqwBin* __restrict__ bins, double* __restrict__ A00, double* __restrict__ S) {
    const int l = blockIdx.x * blockDim.x + threadIdx.x;
    if (l >= n_ell) return;
    double a00 = 0.0, ss = 0.0;
    for (int i = 0; i < n_bin; ++i) {
        const MqwBin b = bins[i];
        const double c = cgg[(size_t)i * cgg_stride + l];
        const double bpNp = b.bp * b.Np, bsNs = b.bs * b.Ns;
        a00 += c * (bpNp * bpNp) + b.wp;
        const double A0i = bpNp * bsNs * c + b.wps;
        const double Aii = bsNs * bsNs * c + b.ws;
        ss += A0i * A0i / Aii;
    }
    A00[l] = a00;
    S[l] = a00 / (a00 - ss);
}

// mqw.py:97-104: block (i, j) sums over l; block (i, n_bin) does the Schur-Limber diagonal of mqw.py:83-87.
__global__ void __launch_bounds__(RED_T)
mqw_fisher_kernel(int n_ell, int n_bin, size_t cgg_stride, const double* __restrict__ cgg, const MqwBin* __restrict__ bins,
                  const double* __restrict__ w /*(2l+1) or 0 beyond an l cut*/, double fsky, const double* __restrict__ A00,
                  const double* __restrict__ S, double* __restrict__ F /*[n_bin][n_bin]*/, double* __restrict__ Fsl /*[n_bin]*/,
                  double* __restrict__ Bd /*NULL, or [n_bin]: cadence flavour -- the diagonal base sum is returned apart*/) {
    __shared__ double sm[RED_T];
    const int i = blockIdx.y, j = blockIdx.x;
    const MqwBin bi = bins[i];
    double acc[1] = {0.0};
    if (j == n_bin) {
        for (int l = threadIdx.x; l < n_ell; l += RED_T) {
            const double c = cgg[(size_t)i * cgg_stride + l];
            const double num = bi.bp * bi.bs * c;
            acc[0] += w[l] * (num * num / A00[l]);
        }
        block_sum<1>(acc, sm);
        if (threadIdx.x == 0) Fsl[i] = bi.Ns * fsky * acc[0];
        return;
    }
    if (j == n_bin + 1) {          // cadence/cadence.py:156: B_i = sum_l (2l+1) S A0i_j^2 / Aii / A00 (no fsky)
        for (int l = threadIdx.x; l < n_ell; l += RED_T) {
            const double ci = cgg[(size_t)i * cgg_stride + l];
            const double bsNs_i = bi.bs * bi.Ns;
            const double Aii = bsNs_i * bsNs_i * ci + bi.ws, A0i_j = bi.bp * bsNs_i * ci;
            acc[0] += w[l] * (S[l] * A0i_j * A0i_j / Aii / A00[l]);
        }
        block_sum<1>(acc, sm);
        if (threadIdx.x == 0) Bd[i] = acc[0];
        return;
    }
    const MqwBin bj = bins[j];
    for (int l = threadIdx.x; l < n_ell; l += RED_T) {
        const double ci = cgg[(size_t)i * cgg_stride + l], cj = cgg[(size_t)j * cgg_stride + l];
        const double a00 = A00[l], s = S[l];
        const double bsNs_i = bi.bs * bi.Ns, bsNs_j = bj.bs * bj.Ns;
        const double Aii = bsNs_i * bsNs_i * ci + bi.ws, Ajj = bsNs_j * bsNs_j * cj + bj.ws;
        const double A0i = bi.bp * bi.Np * bsNs_i * ci + bi.wps, A0j = bj.bp * bj.Np * bsNs_j * cj + bj.wps;
        const double A0i_j = bi.bp * bsNs_i * ci, A0j_j = bj.bp * bsNs_j * cj;
        const double ri = A0i / sqrt(a00 * Aii), rj = A0j / sqrt(a00 * Ajj);
        double t = 2.0 * s * s * ri * rj * sqrt(1.0 / Aii / Ajj) * A0i_j * A0j_j / a00;
        if (i == j && !Bd) t += s * A0i_j * A0i_j / Aii / a00;
        acc[0] += w[l] * t;
    }
    block_sum<1>(acc, sm);
    if (threadIdx.x == 0) F[i * n_bin + j] = fsky * acc[0];
}
}  // namespace lbg

using namespace lbg;

extern "C" int lbg_snr(lbg_ctx* c, int n_ell, const double* ell, const double* cl, const double* var, const double* nmodes,
                       double lmax, double* out_snr) {
    LBG_REQUIRE(c && out_snr && n_ell >= 0, "lbg_snr: bad argument");
    LBG_REQUIRE(n_ell == 0 || (ell && cl && var && nmodes), "lbg_snr: NULL array");
    if (n_ell == 0) { *out_snr = 0.0; return LBG_OK; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t nb = (size_t)n_ell * 8;
    if (c->s0.reserve(4 * nb + 8)) return LBG_ENOMEM;
    double* d = c->s0.as<double>();
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(d, ell, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + n_ell, cl, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + 2 * (size_t)n_ell, var, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + 3 * (size_t)n_ell, nmodes, nb, cudaMemcpyHostToDevice, st));
    KernelTimer t(c);
    snr_kernel<<<1, RED_T, 0, st>>>(n_ell, d, d + n_ell, d + 2 * (size_t)n_ell, d + 3 * (size_t)n_ell, lmax, d + 4 * (size_t)n_ell);
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_snr, d + 4 * (size_t)n_ell, 8, cudaMemcpyDeviceToHost, st));
    return t.finish();
}

extern "C" int lbg_snr_grid(lbg_ctx* c, int n_ell, const double* ell, const double* nmodes, const double* ckk, const double* cgg,
                            const double* ckg, int n_exp, const double* nkk, const double* fsky, int n_rho, const double* rho,
                            int n_dil, const double* ngg, double lmax, double* out) {
    LBG_REQUIRE(c && out, "lbg_snr_grid: NULL argument");
    LBG_REQUIRE(n_ell >= 0 && n_exp >= 1 && n_exp <= 65535 && n_rho >= 1 && n_rho <= 65535 && n_dil >= 1,
                "lbg_snr_grid: bad sizes n_ell=%d n_exp=%d n_rho=%d n_dil=%d", n_ell, n_exp, n_rho, n_dil);
    LBG_REQUIRE(fsky && ngg && (n_ell == 0 || (ell && nmodes && ckk && cgg && ckg && nkk && rho)), "lbg_snr_grid: NULL argument");
    for (int e = 0; e < n_exp; ++e) LBG_REQUIRE(fsky[e] > 0, "lbg_snr_grid: fsky[%d] must be positive", e);
    const size_t no = (size_t)n_exp * n_rho * n_dil;
    if (n_ell == 0) { for (size_t i = 0; i < no; ++i) out[i] = 0.0; return LBG_OK; }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    // one pinned staging block: [ell | nmodes | ckk | cgg | ckg | nkk | rho | fsky | ngg]
    const size_t ne = (size_t)n_ell, n_in = 5 * ne + (size_t)n_exp * ne + (size_t)n_rho * ne + n_exp + n_dil;
    if (c->s0.reserve(n_in * 8) || c->pin_in.reserve(n_in * 8) || c->s4.reserve(no * 8) || c->pin_out.reserve(no * 8)) return LBG_ENOMEM;
    double* h = c->pin_in.as<double>();
    size_t o = 0;
    auto put = [&](const double* src, size_t n) { size_t at = o; memcpy(h + o, src, n * 8); o += n; return at; };
    const size_t o_ell = put(ell, ne), o_nm = put(nmodes, ne), o_kk = put(ckk, ne), o_gg = put(cgg, ne), o_kg = put(ckg, ne);
    const size_t o_nkk = put(nkk, (size_t)n_exp * ne), o_rho = put(rho, (size_t)n_rho * ne), o_fs = put(fsky, n_exp), o_ng = put(ngg, n_dil);
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, h, o * 8, cudaMemcpyHostToDevice, st));
    const double* d = c->s0.as<double>();
    KernelTimer t(c);
    snr_grid_kernel<<<dim3(n_dil, n_rho, n_exp), RED_T, 0, st>>>(n_ell, d + o_ell, d + o_nm, d + o_kk, d + o_gg, d + o_kg, d + o_nkk, d + o_fs,
                                                                 d + o_rho, d + o_ng, lmax, c->s4.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(c->pin_out.p, c->s4.p, no * 8, cudaMemcpyDeviceToHost, st));
    const int rc = t.finish();
    if (rc == LBG_OK) memcpy(out, c->pin_out.p, no * 8);
    return rc;
}

extern "C" int lbg_gaussian_cov(lbg_ctx* c, int n_ell, const double* ell, int n_f, const double* tot, int n_spec,
                                const int* spec_a, const int* spec_b, double fsky, double* out_cov) {
    LBG_REQUIRE(c && ell && tot && spec_a && spec_b && out_cov, "lbg_gaussian_cov: NULL argument");
    LBG_REQUIRE(n_ell >= 0 && n_f > 0 && n_spec > 0 && fsky > 0, "lbg_gaussian_cov: bad sizes");
    for (int i = 0; i < n_spec; ++i)
        LBG_REQUIRE(spec_a[i] >= 0 && spec_a[i] < n_f && spec_b[i] >= 0 && spec_b[i] < n_f, "lbg_gaussian_cov: spectrum %d names a field outside 0..%d", i, n_f - 1);
    if (n_ell == 0) return LBG_OK;
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t nt = (size_t)n_ell * n_f * n_f, nc = (size_t)n_ell * n_spec * n_spec;
    if (c->s0.reserve((size_t)n_ell * 8) || c->s1.reserve(nt * 8) || c->s2.reserve((size_t)2 * n_spec * 4) || c->s3.reserve(nc * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, ell, (size_t)n_ell * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, tot, nt * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s2.p, spec_a, (size_t)n_spec * 4, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s2.as<int>() + n_spec, spec_b, (size_t)n_spec * 4, cudaMemcpyHostToDevice, st));
    KernelTimer t(c);
    gaussian_cov_kernel<<<(unsigned)((nc + 127) / 128), 128, 0, st>>>(n_ell, n_f, n_spec, c->s0.as<double>(), c->s1.as<double>(),
                                                                     c->s2.as<int>(), c->s2.as<int>() + n_spec, fsky, c->s3.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_cov, c->s3.p, nc * 8, cudaMemcpyDeviceToHost, st));
    return t.finish();
}

extern "C" int lbg_fisher_accum(lbg_ctx* c, int n_ell, int n_spec, int n_par, const double* cov, const double* dD,
                                const double* dDelta, const double* mode_weight, double* out_F, double* out_b,
                                double* out_per_ell) {
    LBG_REQUIRE(c && out_F, "lbg_fisher_accum: NULL argument");
    LBG_REQUIRE(n_ell >= 0 && n_spec >= 1 && n_spec <= 32, "lbg_fisher_accum: n_spec=%d outside [1, 32]", n_spec);
    LBG_REQUIRE(n_par >= 1 && n_par <= MAX_PAR, "lbg_fisher_accum: n_par=%d outside [1, %d]", n_par, MAX_PAR);
    LBG_REQUIRE(n_ell == 0 || (cov && dD && mode_weight), "lbg_fisher_accum: NULL array");
    LBG_REQUIRE(!out_b || dDelta, "lbg_fisher_accum: out_b requested without dDelta");
    const int nq = n_par * n_par + n_par;
    if (n_ell == 0) {
        for (int e = 0; e < n_par * n_par; ++e) out_F[e] = 0.0;
        if (out_b) for (int e = 0; e < n_par; ++e) out_b[e] = 0.0;
        return LBG_OK;
    }
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t ncov = (size_t)n_ell * n_spec * n_spec, nd = (size_t)n_ell * n_par * n_spec, ndl = (size_t)n_ell * n_spec;
    if (c->s0.reserve(ncov * 8) || c->s1.reserve(nd * 8) || c->s2.reserve(ndl * 8) || c->s3.reserve((size_t)n_ell * 8) ||
        c->s4.reserve((size_t)n_ell * nq * 8) || c->s5.reserve((size_t)nq * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, cov, ncov * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, dD, nd * 8, cudaMemcpyHostToDevice, st));
    if (dDelta) LBG_CUDA(cudaMemcpyAsync(c->s2.p, dDelta, ndl * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s3.p, mode_weight, (size_t)n_ell * 8, cudaMemcpyHostToDevice, st));
    const double* dd = dDelta ? c->s2.as<double>() : nullptr;
    const int gx = (n_ell + 63) / 64;
    KernelTimer t(c);
    if (n_spec <= 4)       quadform_kernel<4><<<gx, 64, 0, st>>>(n_ell, n_spec, n_par, c->s0.as<double>(), c->s1.as<double>(), dd, c->s4.as<double>());
    else if (n_spec <= 8)  quadform_kernel<8><<<gx, 64, 0, st>>>(n_ell, n_spec, n_par, c->s0.as<double>(), c->s1.as<double>(), dd, c->s4.as<double>());
    else if (n_spec <= 16) quadform_kernel<16><<<gx, 64, 0, st>>>(n_ell, n_spec, n_par, c->s0.as<double>(), c->s1.as<double>(), dd, c->s4.as<double>());
    else                   quadform_kernel<32><<<gx, 64, 0, st>>>(n_ell, n_spec, n_par, c->s0.as<double>(), c->s1.as<double>(), dd, c->s4.as<double>());
    t.launched();
    weighted_sum_kernel<<<nq, RED_T, 0, st>>>(n_ell, nq, c->s3.as<double>(), c->s4.as<double>(), c->s5.as<double>());
    t.launched();
    std::vector<double> host(nq);
    LBG_CUDA(cudaMemcpyAsync(host.data(), c->s5.p, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    std::vector<double> per;
    if (out_per_ell) {
        per.resize((size_t)n_ell * nq);
        LBG_CUDA(cudaMemcpyAsync(per.data(), c->s4.p, per.size() * 8, cudaMemcpyDeviceToHost, st));
    }
    int rc = t.finish();
    if (rc) return rc;
    memcpy(out_F, host.data(), (size_t)n_par * n_par * 8);
    if (out_b) memcpy(out_b, host.data() + n_par * n_par, (size_t)n_par * 8);
    if (out_per_ell)
        for (int l = 0; l < n_ell; ++l) memcpy(out_per_ell + (size_t)l * n_par * n_par, per.data() + (size_t)l * nq, (size_t)n_par * n_par * 8);
    return LBG_OK;
}

// Field-level Fisher matrix for n_f correlated Gaussian fields (kappa and every tomographic galaxy bin at once):
//   Q_l[p][q] = 1/2 Tr[ C_l^-1 d_p C_l  C_l^-1 d_q C_l ],   C_l[a][b] = (C + N)^{ab}(l),
// algebraically the interloper.py:223-253 sum dD^T Cov^-1 dD with the Gaussian covariance of interloper.py:42-63 taken over ALL
// auto and cross spectra of the fields (the (2l+1) fsky factor is part of the caller's mode weight), but without ever forming the
// n_spec x n_spec covariance: 121 fields (6 tracers x 20 bins + kappa, BASELINE configs[3]) would need a 7381 x 7381 inverse per l.
// One CTA per multipole (persistent, strided), everything in shared memory:
//   1. C_l scattered from the spectra list, inverted in place (Gauss-Jordan without pivoting: C_l is symmetric positive definite;
//      warp <-> row, lane <-> column, two barriers per pivot);
//   2. for every (i, j): a_p = (W d_pC)[i][j] from the CSC column j of the pattern and b_q = (d_qC W)[i][j] from column i
//      (the pattern is symmetric), Q_pq += a_p b_q -- nothing is stored.  Columns much denser than the rest (the kappa column of
//      the arrowhead pattern: n_f entries against 2) would serialise the warps that meet them, so up to FF_HEAVY such columns are
//      evaluated once per multipole by all threads into shared memory (W d_pC)[.][h] and read from there;
//   3. fixed-order block reduction of the n_par (n_par + 1) / 2 sums.
constexpr int FF_T = 512;
constexpr int FF_HEAVY = 2;

// One Gauss-Jordan sweep over the rows a warp holds, pivot column in column group KU: multiplier from the owning lane, one FMA per element.
// (The pivot row passes through here too and is overwritten by the caller.)
template <int NCG, int RPW, int KU>
__device__ __forceinline__ void ff_sweep(double (&m)[RPW][NCG], const double (&y)[NCG], double d, int kl, int lane) {
    const bool own = lane == kl;
#pragma unroll
    for (int r = 0; r < RPW; ++r) {
        const double f = __shfl_sync(0xffffffffu, m[r][KU], kl) * d;
#pragma unroll
        for (int u = 0; u < NCG; ++u) m[r][u] = m[r][u] - f * y[u];
        if (own) m[r][KU] = -f;
    }
}

template <bool DV_SM, int NCG /*32-column groups per row: row stride S = 32 NCG >= n_f, padding kept zero*/>
__global__ void __launch_bounds__(FF_T)
field_fisher_kernel(int n_ell, int nf, int n_spec, int np, int n_heavy, const int* __restrict__ sa, const int* __restrict__ sb,
                    const int* __restrict__ colptr, const int* __restrict__ rowidx, const int* __restrict__ specidx,
                    const int* __restrict__ heavy /*[nf]: slot of a heavy column or -1*/,
                    const double* __restrict__ tot /*[n_spec][n_ell]*/, const double* __restrict__ dC /*[np][n_spec][n_ell]*/,
                    double* __restrict__ q /*[n_ell][np*np]*/) {
    extern __shared__ double ff_sm[];
    constexpr int S = 32 * NCG, RPW = 2 * NCG;               // rows per warp: 16 warps x RPW >= S >= n_f
    double* W = ff_sm;                                       // [nf][S]
    double* red = W + (size_t)nf * S;                        // [FF_T / 32][36]
    double* pbuf = red + (FF_T / 32) * 36;                   // [2][S] pivot row, double-buffered
    double* mh = pbuf + 2 * S;                     // [n_heavy][np][nf]
    double* dval = mh + (size_t)n_heavy * np * nf;           // [np][n_spec] (DV_SM)
    const int n2 = nf * S, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    __shared__ int hcol[FF_HEAVY];
    for (int j = tid; j < nf; j += FF_T) if (heavy[j] >= 0) hcol[heavy[j]] = j;
    for (int l = blockIdx.x; l < n_ell; l += gridDim.x) {
        for (int e = tid; e < n2; e += FF_T) W[e] = 0.0;
        if (DV_SM) for (int e = tid; e < np * n_spec; e += FF_T) dval[e] = dC[(size_t)e * n_ell + l];
        __syncthreads();
        for (int s = tid; s < n_spec; s += FF_T) {
            const double v = tot[(size_t)s * n_ell + l];
            W[sa[s] * S + sb[s]] = v;
            W[sb[s] * S + sa[s]] = v;
        }
        __syncthreads();
        // Gauss-Jordan with the matrix in REGISTERS: warp w holds rows w, w + 16, .. (slot r), lane <-> columns lane, lane + 32, ..
        // (zero padding stays zero).  Per pivot: the owning warp publishes row k through a double-buffered shared row (one barrier per
        // pivot), every row takes its multiplier from the lane that owns column k by shuffle, one FMA per element.  Shared memory
        // carries 2 rows per pivot instead of the whole matrix twice (the shared-memory version sat at 79 % of the L1 data pipe).
        double m[RPW][NCG];
#pragma unroll
        for (int r = 0; r < RPW; ++r) {
            const int i = warp + (FF_T / 32) * r;
#pragma unroll
            for (int u = 0; u < NCG; ++u) m[r][u] = i < nf ? W[i * S + lane + 32 * u] : 0.0;
        }
        for (int k = 0; k < nf; ++k) {
            double* prow = pbuf + (k & 1) * S;
            const int kw = k & (FF_T / 32 - 1), kr = k / (FF_T / 32), ku = k >> 5, kl = k & 31;
            if (warp == kw) {
#pragma unroll
                for (int r = 0; r < RPW; ++r)
                    if (r == kr) {
#pragma unroll
                        for (int u = 0; u < NCG; ++u) prow[lane + 32 * u] = m[r][u];
                    }
            }
            __syncthreads();
            const double d = 1.0 / prow[k];
            double y[NCG];
#pragma unroll
            for (int u = 0; u < NCG; ++u) y[u] = prow[lane + 32 * u];
            // the column group of k is warp-uniform: one specialised sweep per group keeps every register index static
            switch (ku) {
                case 0: ff_sweep<NCG, RPW, 0>(m, y, d, kl, lane); break;
                case 1: if constexpr (NCG > 1) ff_sweep<NCG, RPW, 1>(m, y, d, kl, lane); break;
                case 2: if constexpr (NCG > 2) ff_sweep<NCG, RPW, 2>(m, y, d, kl, lane); break;
                case 3: if constexpr (NCG > 3) ff_sweep<NCG, RPW, 3>(m, y, d, kl, lane); break;
                default: if constexpr (NCG > 4) ff_sweep<NCG, RPW, 4>(m, y, d, kl, lane); break;
            }
            if (warp == kw) {                                // the pivot row itself: scaled old row, pivot -> 1 / pivot
#pragma unroll
                for (int r = 0; r < RPW; ++r)
                    if (r == kr) {
#pragma unroll
                        for (int u = 0; u < NCG; ++u) m[r][u] = (u == ku && lane == kl) ? d : y[u] * d;
                    }
            }
        }
#pragma unroll
        for (int r = 0; r < RPW; ++r) {
            const int i = warp + (FF_T / 32) * r;
            if (i < nf) {
#pragma unroll
                for (int u = 0; u < NCG; ++u) W[i * S + lane + 32 * u] = m[r][u];
            }
        }
        __syncthreads();
        // dv(p, s): derivative p of spectrum s at this multipole
#define FF_DV(p, s) (DV_SM ? dval[(p) * n_spec + (s)] : dC[((size_t)(p) * n_spec + (s)) * n_ell + l])
        for (int e = tid; e < n_heavy * np * nf; e += FF_T) {            // (W d_pC)[i][h] for the heavy columns h
            const int i = e % nf, p = (e / nf) % np, h = e / (nf * np), col = hcol[h];
            double acc = 0.0;
            for (int t = colptr[col]; t < colptr[col + 1]; ++t) acc += W[rowidx[t] * S + i] * FF_DV(p, specidx[t]);   // W symmetric
            mh[e] = acc;
        }
        __syncthreads();
        double acc[36];
#pragma unroll
        for (int e = 0; e < 36; ++e) acc[e] = 0.0;
        for (int i = warp; i < nf; i += FF_T / 32) {
            const int hi = heavy[i], i0 = colptr[i], i1 = colptr[i + 1];
            for (int j = lane; j < nf; j += 32) {
                const int hj = heavy[j];
                double a[MAX_PAR], b[MAX_PAR];
#pragma unroll
                for (int p = 0; p < MAX_PAR; ++p) a[p] = b[p] = 0.0;
                if (hj >= 0) {
#pragma unroll
                    for (int p = 0; p < MAX_PAR; ++p) if (p < np) a[p] = mh[(hj * np + p) * nf + i];
                } else {
                    for (int t = colptr[j]; t < colptr[j + 1]; ++t) {
                        const double w = W[i * S + rowidx[t]];
                        const int s = specidx[t];
#pragma unroll
                        for (int p = 0; p < MAX_PAR; ++p) if (p < np) a[p] += w * FF_DV(p, s);
                    }
                }
                if (hi >= 0) {
#pragma unroll
                    for (int p = 0; p < MAX_PAR; ++p) if (p < np) b[p] = mh[(hi * np + p) * nf + j];
                } else {
                    for (int t = i0; t < i1; ++t) {
                        const double w = W[rowidx[t] * S + j];
                        const int s = specidx[t];
#pragma unroll
                        for (int p = 0; p < MAX_PAR; ++p) if (p < np) b[p] += w * FF_DV(p, s);
                    }
                }
                int t = 0;
#pragma unroll
                for (int p = 0; p < MAX_PAR; ++p)
#pragma unroll
                    for (int r = p; r < MAX_PAR; ++r, ++t) acc[t] += a[p] * b[r];
            }
        }
#undef FF_DV
#pragma unroll
        for (int e = 0; e < 36; ++e) {
            double v = acc[e];
            for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
            if (lane == 0) red[warp * 36 + e] = v;
        }
        __syncthreads();
        if (tid < np * np) {
            const int p = tid / np, r = tid - p * np, lo = p < r ? p : r, hi = p < r ? r : p;
            const int t = lo * MAX_PAR - lo * (lo - 1) / 2 + (hi - lo);
            double v = 0.0;
            for (int w = 0; w < FF_T / 32; ++w) v += red[w * 36 + t];
            q[(size_t)l * np * np + tid] = 0.5 * v;
        }
        __syncthreads();
    }
}

// dC rows of amplitude-type parameters built on the device: dC[n_dense + a][s][l] = amp[a][s] * sig[s][l]
__global__ void ff_expand_amp_kernel(int n_amp, int n_spec, int n_ell, const double* __restrict__ amp, const double* __restrict__ sig,
                                     double* __restrict__ out) {
    const size_t nt = (size_t)n_spec * n_ell;
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nt) return;
    const int sp = (int)(t / n_ell);
    const double v = sig[t];
    for (int a = 0; a < n_amp; ++a) out[(size_t)a * nt + t] = amp[(size_t)a * n_spec + sp] * v;
}

// n_par = n_dense + n_amp parameters: the first n_dense with explicit derivative rows dC[n_dense][n_spec][n_ell], the others amplitude
// parameters described by amp[n_amp][n_spec] (d C_s / d A_a = amp[a][s] * sig_s) and the signal spectra sig[n_spec][n_ell].
// Inputs taken from spectra sets kept on the device (lbg_limber_keep / lbg_spectra_set_fill) instead of host arrays:
//   tot = fid + noise,   dC_j = (plus_j - minus_j) / two_h_j  (central difference, the same two roundings as numpy's subtract and divide),
//   sig = fid
struct FFSets { int fid, noise, n_diff; int plus[MAX_PAR], minus[MAX_PAR]; double two_h[MAX_PAR]; };
struct FFAsm { const double* fid; const double* noise; const double* plus[MAX_PAR]; const double* minus[MAX_PAR]; double two_h[MAX_PAR]; int n_diff; };

__global__ void ff_assemble_kernel(size_t nt, FFAsm a, double* __restrict__ tot, double* __restrict__ dC) {
    const size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nt) return;
    const double f = a.fid[t];
    tot[t] = a.noise ? __dadd_rn(f, a.noise[t]) : f;
    for (int j = 0; j < a.n_diff; ++j) dC[(size_t)j * nt + t] = __ddiv_rn(__dsub_rn(a.plus[j][t], a.minus[j][t]), a.two_h[j]);
}

static int field_fisher_impl(lbg_ctx* c, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, const double* tot,
                             int n_dense, const double* dC, int n_amp, const double* amp, const double* sig,
                             const double* mode_weight, double* out_F, double* out_per_ell, const FFSets* sets = nullptr) {
    if (sets) n_dense = sets->n_diff;
    const int n_par = n_dense + n_amp;
    LBG_REQUIRE(c && out_F, "lbg_field_fisher: NULL argument");
    LBG_REQUIRE(n_ell >= 0 && n_f >= 1 && n_f <= 160, "lbg_field_fisher: n_f=%d outside [1, 160]", n_f);
    LBG_REQUIRE(n_spec >= 1 && n_spec <= n_f * (n_f + 1) / 2, "lbg_field_fisher: n_spec=%d outside [1, n_f (n_f + 1) / 2]", n_spec);
    LBG_REQUIRE(n_par >= 1 && n_par <= MAX_PAR, "lbg_field_fisher: n_par=%d outside [1, %d]", n_par, MAX_PAR);
    LBG_REQUIRE(spec_a && spec_b, "lbg_field_fisher: NULL spectrum list");
    LBG_REQUIRE(n_dense >= 0 && n_amp >= 0, "lbg_field_fisher: negative parameter count");
    LBG_REQUIRE(n_ell == 0 || sets || (tot && (n_dense == 0 || dC) && (n_amp == 0 || sig)), "lbg_field_fisher: NULL array");
    LBG_REQUIRE(n_ell == 0 || (mode_weight && (n_amp == 0 || amp)), "lbg_field_fisher: NULL array");
    // CSC pattern of the symmetric matrices the spectra list fills; every field needs its auto spectrum (else C_l is singular)
    std::vector<int> cnt(n_f + 1, 0), seen((size_t)n_f * n_f, 0);
    for (int s = 0; s < n_spec; ++s) {
        const int a = spec_a[s], b = spec_b[s];
        LBG_REQUIRE(a >= 0 && a < n_f && b >= 0 && b < n_f, "lbg_field_fisher: spectrum %d names field pair (%d, %d) outside [0, %d)", s, a, b, n_f);
        LBG_REQUIRE(!seen[(size_t)a * n_f + b], "lbg_field_fisher: field pair (%d, %d) listed twice", a, b);
        seen[(size_t)a * n_f + b] = seen[(size_t)b * n_f + a] = 1;
        cnt[b + 1]++;
        if (a != b) cnt[a + 1]++;
    }
    for (int f = 0; f < n_f; ++f) LBG_REQUIRE(seen[(size_t)f * n_f + f], "lbg_field_fisher: field %d has no auto spectrum", f);
    for (int e = 0; e < n_par * n_par; ++e) out_F[e] = 0.0;
    if (n_ell == 0) return LBG_OK;
    // heavy columns: more than 8 entries and more than 4x the mean column; staged only when there are at most FF_HEAVY of them
    const int nnz = 2 * n_spec - n_f;
    std::vector<int> hv(n_f, -1);
    int n_heavy = 0;
    for (int f = 0; f < n_f; ++f)
        if (cnt[f + 1] > 8 && (long long)cnt[f + 1] * n_f > 4LL * nnz) { if (n_heavy < FF_HEAVY) hv[f] = n_heavy; n_heavy++; }
    if (n_heavy > FF_HEAVY) { n_heavy = 0; std::fill(hv.begin(), hv.end(), -1); }
    for (int f = 0; f < n_f; ++f) cnt[f + 1] += cnt[f];
    std::vector<int> pat((size_t)2 * n_spec + 2 * n_f + 1 + 2 * (size_t)nnz), fill(cnt.begin(), cnt.end() - 1);
    int* h_sa = pat.data(); int* h_sb = h_sa + n_spec; int* h_cp = h_sb + n_spec; int* h_hv = h_cp + n_f + 1; int* h_ri = h_hv + n_f;
    int* h_si = h_ri + nnz;
    for (int s = 0; s < n_spec; ++s) {
        const int a = spec_a[s], b = spec_b[s];
        h_sa[s] = a; h_sb[s] = b;
        h_ri[fill[b]] = a; h_si[fill[b]++] = s;
        if (a != b) { h_ri[fill[a]] = b; h_si[fill[a]++] = s; }
    }
    for (int f = 0; f <= n_f; ++f) h_cp[f] = cnt[f];
    for (int f = 0; f < n_f; ++f) h_hv[f] = hv[f];
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const int ncg = (n_f + 31) / 32;
    const size_t n2 = (size_t)n_f * 32 * ncg;
    size_t base = (n2 + (FF_T / 32) * 36 + 64 * ncg + (size_t)n_heavy * n_par * n_f) * 8;
    if (base > 226u * 1024u) {                 // no room to stage the dense columns next to a 160 x 160 matrix: evaluate them in place
        n_heavy = 0;
        for (int f = 0; f < n_f; ++f) h_hv[f] = -1;
        base = (n2 + (FF_T / 32) * 36 + 64 * ncg) * 8;
    }
    const size_t dvb = (size_t)n_par * n_spec * 8;
    const bool dv_sm = base + dvb <= 220u * 1024u;
    const size_t smem = base + (dv_sm ? dvb : 0);
    int sms = 148;
    LBG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, c->device));
    const int gx = n_ell < sms ? n_ell : sms;     // one CTA per SM: 512 threads x 128 registers fill the register file
    const int nq = n_par * n_par;
    const size_t nt = (size_t)n_spec * n_ell, nd = nt * n_par;
    if (c->s0.reserve(nt * 8) || c->s1.reserve((nd + (n_amp ? nt : 0)) * 8) || c->s2.reserve(pat.size() * 4) ||
        c->s3.reserve(((size_t)n_ell + (size_t)n_amp * n_spec) * 8) ||
        c->s4.reserve((size_t)n_ell * nq * 8) || c->s5.reserve((size_t)nq * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    typedef void (*ff_kernel_t)(int, int, int, int, int, const int*, const int*, const int*, const int*, const int*, const int*, const double*,
                                const double*, double*);
    static const ff_kernel_t kerns[2][5] = {
        {field_fisher_kernel<false, 1>, field_fisher_kernel<false, 2>, field_fisher_kernel<false, 3>, field_fisher_kernel<false, 4>, field_fisher_kernel<false, 5>},
        {field_fisher_kernel<true, 1>, field_fisher_kernel<true, 2>, field_fisher_kernel<true, 3>, field_fisher_kernel<true, 4>, field_fisher_kernel<true, 5>}};
    const ff_kernel_t kern = kerns[dv_sm ? 1 : 0][ncg - 1];
    LBG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    FFAsm fa{};
    if (sets) {
        auto set_ptr = [&](int s, const double** out) -> int {
            LBG_REQUIRE(s >= 0 && s < lbg_ctx::SPEC_SETS, "lbg_field_fisher_sets: set %d outside [0, %d)", s, lbg_ctx::SPEC_SETS);
            if (c->set_rows[s] != n_spec || c->set_n_ell[s] != n_ell) {
                set_error("lbg_field_fisher_sets: set %d holds %d rows x %d multipoles, the call names %d spectra x %d multipoles", s,
                          c->set_rows[s], c->set_n_ell[s], n_spec, n_ell);
                return LBG_ESTATE;
            }
            *out = c->spec_set[s].as<double>();
            return LBG_OK;
        };
        if (int rc = set_ptr(sets->fid, &fa.fid)) return rc;
        if (sets->noise >= 0) { if (int rc = set_ptr(sets->noise, &fa.noise)) return rc; }
        fa.n_diff = sets->n_diff;
        for (int j = 0; j < sets->n_diff; ++j) {
            if (int rc = set_ptr(sets->plus[j], &fa.plus[j])) return rc;
            if (int rc = set_ptr(sets->minus[j], &fa.minus[j])) return rc;
            fa.two_h[j] = sets->two_h[j];
        }
    } else {
        LBG_CUDA(cudaMemcpyAsync(c->s0.p, tot, nt * 8, cudaMemcpyHostToDevice, st));
        if (n_dense) LBG_CUDA(cudaMemcpyAsync(c->s1.p, dC, nt * n_dense * 8, cudaMemcpyHostToDevice, st));
        if (n_amp) LBG_CUDA(cudaMemcpyAsync(c->s1.as<double>() + nd, sig, nt * 8, cudaMemcpyHostToDevice, st));
    }
    LBG_CUDA(cudaMemcpyAsync(c->s2.p, pat.data(), pat.size() * 4, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s3.p, mode_weight, (size_t)n_ell * 8, cudaMemcpyHostToDevice, st));
    if (n_amp) LBG_CUDA(cudaMemcpyAsync(c->s3.as<double>() + n_ell, amp, (size_t)n_amp * n_spec * 8, cudaMemcpyHostToDevice, st));
    const int* d_pat = c->s2.as<int>();
    KernelTimer t(c);
    if (sets) {
        ff_assemble_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(nt, fa, c->s0.as<double>(), c->s1.as<double>());
        t.launched();
    }
    if (n_amp) {
        ff_expand_amp_kernel<<<(unsigned)((nt + 255) / 256), 256, 0, st>>>(n_amp, n_spec, n_ell, c->s3.as<double>() + n_ell,
                                                                           sets ? fa.fid : c->s1.as<double>() + nd, c->s1.as<double>() + nt * n_dense);
        t.launched();
    }
    kern<<<gx, FF_T, smem, st>>>(n_ell, n_f, n_spec, n_par, n_heavy, d_pat, d_pat + n_spec, d_pat + 2 * n_spec,
                                 d_pat + 2 * n_spec + 2 * n_f + 1, d_pat + 2 * n_spec + 2 * n_f + 1 + nnz, d_pat + 2 * n_spec + n_f + 1,
                                 c->s0.as<double>(), c->s1.as<double>(), c->s4.as<double>());
    t.launched();
    weighted_sum_kernel<<<nq, RED_T, 0, st>>>(n_ell, nq, c->s3.as<double>(), c->s4.as<double>(), c->s5.as<double>());
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_F, c->s5.p, (size_t)nq * 8, cudaMemcpyDeviceToHost, st));
    if (out_per_ell) LBG_CUDA(cudaMemcpyAsync(out_per_ell, c->s4.p, (size_t)n_ell * nq * 8, cudaMemcpyDeviceToHost, st));
    return t.finish();
}

extern "C" int lbg_field_fisher(lbg_ctx* c, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, const double* tot,
                                int n_par, const double* dC, const double* mode_weight, double* out_F, double* out_per_ell) {
    return field_fisher_impl(c, n_ell, n_f, n_spec, spec_a, spec_b, tot, n_par, dC, 0, nullptr, nullptr, mode_weight, out_F, out_per_ell);
}

extern "C" int lbg_field_fisher_amp(lbg_ctx* c, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, const double* tot,
                                    int n_dense, const double* dC, int n_amp, const double* amp, const double* sig,
                                    const double* mode_weight, double* out_F, double* out_per_ell) {
    return field_fisher_impl(c, n_ell, n_f, n_spec, spec_a, spec_b, tot, n_dense, dC, n_amp, amp, sig, mode_weight, out_F, out_per_ell);
}

extern "C" int lbg_field_fisher_sets(lbg_ctx* c, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, int set_fid,
                                     int set_noise, int n_diff, const int* set_plus, const int* set_minus, const double* two_h,
                                     int n_amp, const double* amp, const double* mode_weight, double* out_F, double* out_per_ell) {
    LBG_REQUIRE(n_diff >= 0 && n_diff <= MAX_PAR, "lbg_field_fisher_sets: n_diff=%d outside [0, %d]", n_diff, MAX_PAR);
    LBG_REQUIRE(n_diff == 0 || (set_plus && set_minus && two_h), "lbg_field_fisher_sets: NULL argument");
    FFSets fs{};
    fs.fid = set_fid; fs.noise = set_noise; fs.n_diff = n_diff;
    for (int j = 0; j < n_diff; ++j) {
        LBG_REQUIRE(two_h[j] != 0.0, "lbg_field_fisher_sets: zero step for difference %d", j);
        fs.plus[j] = set_plus[j]; fs.minus[j] = set_minus[j]; fs.two_h[j] = two_h[j];
    }
    return field_fisher_impl(c, n_ell, n_f, n_spec, spec_a, spec_b, nullptr, n_diff, nullptr, n_amp, amp, nullptr, mode_weight, out_F, out_per_ell, &fs);
}

static int mqw_fisher_impl(lbg_ctx* c, int n_ell, int n_bin, const double* cgg, int cgg_per_bin, const double* bins,
                           const double* mode_weight, double fsky, double* out_F, double* out_Fsl, double* out_S, bool cadence) {
    LBG_REQUIRE(c && cgg && bins && mode_weight && out_F, "lbg_mqw_fisher / lbg_cadence_fisher: NULL argument");
    LBG_REQUIRE(n_ell > 0 && n_bin > 0 && n_bin <= 1024, "lbg_mqw_fisher: bad sizes n_ell=%d n_bin=%d", n_ell, n_bin);
    std::lock_guard<std::mutex> g(c->mu);
    LBG_CUDA(cudaSetDevice(c->device));
    const size_t ncg = (size_t)(cgg_per_bin ? n_bin : 1) * n_ell;
    if (c->s0.reserve(ncg * 8) || c->s1.reserve((size_t)n_bin * sizeof(MqwBin)) || c->s2.reserve((size_t)n_ell * 8) ||
        c->s3.reserve((size_t)2 * n_ell * 8) || c->s4.reserve((size_t)(n_bin + 2) * n_bin * 8)) return LBG_ENOMEM;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(c->s0.p, cgg, ncg * 8, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s1.p, bins, (size_t)n_bin * sizeof(MqwBin), cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(c->s2.p, mode_weight, (size_t)n_ell * 8, cudaMemcpyHostToDevice, st));
    double* A00 = c->s3.as<double>();
    double* S = A00 + n_ell;
    double* F = c->s4.as<double>();
    double* Fsl = F + (size_t)n_bin * n_bin;
    double* Bd = cadence ? Fsl + n_bin : nullptr;
    const size_t stride = cgg_per_bin ? (size_t)n_ell : 0;
    KernelTimer t(c);
    mqw_schur_kernel<<<(n_ell + 127) / 128, 128, 0, st>>>(n_ell, n_bin, stride, c->s0.as<double>(), c->s1.as<MqwBin>(), A00, S);
    t.launched();
    mqw_fisher_kernel<<<dim3(n_bin + (cadence ? 2 : 1), n_bin), RED_T, 0, st>>>(n_ell, n_bin, stride, c->s0.as<double>(), c->s1.as<MqwBin>(),
                                                               c->s2.as<double>(), fsky, A00, S, F, Fsl, Bd);
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_F, F, (size_t)n_bin * n_bin * 8, cudaMemcpyDeviceToHost, st));
    if (out_Fsl) LBG_CUDA(cudaMemcpyAsync(out_Fsl, Fsl, (size_t)n_bin * 8, cudaMemcpyDeviceToHost, st));
    if (out_S) LBG_CUDA(cudaMemcpyAsync(out_S, S, (size_t)n_ell * 8, cudaMemcpyDeviceToHost, st));
    std::vector<double> B(cadence ? n_bin : 0);
    if (cadence) LBG_CUDA(cudaMemcpyAsync(B.data(), Bd, (size_t)n_bin * 8, cudaMemcpyDeviceToHost, st));
    int rc = t.finish();
    if (rc || !cadence) return rc;
    // cadence/cadence.py:155-165 as written: `Interim` starts at the base term WITHOUT fsky and keeps accumulating the j terms,
    // and each Fisher[i, j] receives its running sum:  F_ij = [i == j] fsky B_i + B_i + sum_{j' <= j} T_ij'  (out_F holds T)
    for (int i = 0; i < n_bin; ++i) {
        double run = B[i];
        for (int j = 0; j < n_bin; ++j) {
            run += out_F[(size_t)i * n_bin + j];
            out_F[(size_t)i * n_bin + j] = (i == j ? fsky * B[i] : 0.0) + run;
        }
    }
    return LBG_OK;
}

extern "C" int lbg_mqw_fisher(lbg_ctx* c, int n_ell, int n_bin, const double* cgg, int cgg_per_bin, const double* bins,
                              const double* mode_weight, double fsky, double* out_F, double* out_Fsl, double* out_S) {
    return mqw_fisher_impl(c, n_ell, n_bin, cgg, cgg_per_bin, bins, mode_weight, fsky, out_F, out_Fsl, out_S, false);
}

extern "C" int lbg_cadence_fisher(lbg_ctx* c, int n_ell, int n_bin, const double* cgg, const double* bins,
                                  const double* mode_weight, double fsky, double* out_F, double* out_Fsl, double* out_S) {
    return mqw_fisher_impl(c, n_ell, n_bin, cgg, 1, bins, mode_weight, fsky, out_F, out_Fsl, out_S, true);
}

// ---- lbg_snr without any allocation ---------------------------------------------------------------------------------------------
// Some callers cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  lbg_snr stages
// its four input vectors and its result in the context's grow-only scratch buffer (cudaMalloc on first use / growth); this variant
// runs the SAME kernel with that scratch in a file-scope device array of fixed size instead: it calls no cudaMalloc, cudaFree or any
// allocating routine.  Every GPU that loads the library holds the array (LBG_SNR_NOALLOC_MAX x 4 doubles + one line = 512 KiB + 256 B)
// until the process exits.  Every call on a device shares the array, and it keeps the last call's contents: the result slot is
// cleared before each launch, and calls that overlap on one GPU (two contexts / lanes of a device, or two threads) race -- they
// are serialised here by a process-wide mutex, callers outside this library's knowledge must not overlap them themselves.
#define LBG_SNR_NOALLOC_MAX 16384
static __device__ __align__(256) double g_snr_scratch[4 * LBG_SNR_NOALLOC_MAX + 32];   // [ell | cl | var | nmodes | result (own 256 B line)]

namespace {
std::mutex g_snr_noalloc_mu;
double* g_snr_scratch_ptr[64];          // device address of g_snr_scratch per device (in host code the array's NAME is only a stand-in)
}

extern "C" int lbg_snr_noalloc(lbg_ctx* c, int n_ell, const double* ell, const double* cl, const double* var, const double* nmodes,
                               double lmax, double* out_snr) {
    LBG_REQUIRE(c && out_snr && n_ell >= 0, "lbg_snr_noalloc: bad argument");
    LBG_REQUIRE(n_ell == 0 || (ell && cl && var && nmodes), "lbg_snr_noalloc: NULL array");
    LBG_REQUIRE(n_ell <= LBG_SNR_NOALLOC_MAX, "lbg_snr_noalloc: n_ell=%d exceeds the fixed scratch of %d multipoles (use lbg_snr, which allocates)",
                n_ell, LBG_SNR_NOALLOC_MAX);
    LBG_REQUIRE(c->device >= 0 && c->device < 64, "lbg_snr_noalloc: device %d outside the 64 devices this entry point keeps an address for", c->device);
    if (n_ell == 0) { *out_snr = 0.0; return LBG_OK; }
    std::lock_guard<std::mutex> g(c->mu);
    std::lock_guard<std::mutex> gs(g_snr_noalloc_mu);          // the array is shared by every context of the device
    LBG_CUDA(cudaSetDevice(c->device));
    double*& d = g_snr_scratch_ptr[c->device];
    if (!d) {                                                  // set-up, once per device, with that device current
        double* p = nullptr;
        LBG_CUDA(cudaGetSymbolAddress((void**)&p, g_snr_scratch));
        d = p;
    }
    const size_t nb = (size_t)n_ell * 8, N = LBG_SNR_NOALLOC_MAX;
    double* d_out = d + 4 * N;
    cudaStream_t st = c->stream;
    LBG_CUDA(cudaMemcpyAsync(d, ell, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + N, cl, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + 2 * N, var, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemcpyAsync(d + 3 * N, nmodes, nb, cudaMemcpyHostToDevice, st));
    LBG_CUDA(cudaMemsetAsync(d_out, 0, 8, st));                // the slot keeps the previous call's result otherwise
    KernelTimer t(c);
    snr_kernel<<<1, RED_T, 0, st>>>(n_ell, d, d + N, d + 2 * N, d + 3 * N, lmax, d_out);
    t.launched();
    LBG_CUDA(cudaMemcpyAsync(out_snr, d_out, 8, cudaMemcpyDeviceToHost, st));
    return t.finish();
}
