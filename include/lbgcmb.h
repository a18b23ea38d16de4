/* lbgcmb.h -- C ABI of liblbgcmb.so, the B200 (sm_100a) implementation of the LBGCMB forecast hot path.
 *
 * The reference (michaelJwilson/LBGCMB) has no FFI: its boundary is a set of plain Python functions
 * (SURVEY.md 8(b)).  Each entry point below names the reference function whose arithmetic it replaces;
 * the Python mirrors in lbgcmb_b200/ (same names / argument meaning as the reference) are thin ctypes
 * callers of these.  All pointers are HOST pointers to C-contiguous arrays owned by the caller; the
 * library copies host<->device itself.  Every function returns 0 on success or a negative LBG_E* code;
 * lbg_last_error() gives the message of the last failure on the calling thread.
 * There is no CPU fallback: without a CUDA device lbg_ctx_create() fails.
 */
#ifndef LBGCMB_H
#define LBGCMB_H

#ifdef __cplusplus
extern "C" {
#endif

#define LBG_OK          0
#define LBG_EINVAL     -1   /* bad argument                     */
#define LBG_ECUDA      -2   /* CUDA runtime error               */
#define LBG_ESTATE     -3   /* required table not uploaded yet  */
#define LBG_ENOMEM     -4

/* estimator bits for `terms_mask` (cmb/Nkk.py:31 `terms`) */
#define LBG_TT  1u
#define LBG_TE  2u
#define LBG_EE  4u
#define LBG_EB  8u
#define LBG_BB 16u

typedef struct lbg_ctx lbg_ctx;   /* opaque: device, stream, uploaded tables, workspaces */

/* ---- context ------------------------------------------------------------------------------------------- */
int         lbg_ctx_create(int device, lbg_ctx** out);
/* A further context on the same GPU (own stream, scratch and mutex) whose calls may overlap those of lane 0: lane 0 streams have
 * the device's highest priority, lanes >= 1 the lowest, so a long N0 launch on lane 1 yields SM slots to the short Limber /
 * covariance kernels of lane 0 (lbgcmb_b200/Nkk.py n0_terms_async).  lbg_ctx_create(d, out) == lbg_ctx_create_lane(d, 0, out). */
int         lbg_ctx_create_lane(int device, int lane, lbg_ctx** out);
int         lbg_ctx_destroy(lbg_ctx* ctx);
const char* lbg_last_error(void);
const char* lbg_version(void);
/* Device time (ms, CUDA events on the context's stream) spent in the kernels of the last compute call,
 * and the number of kernel launches that call made. */
int         lbg_last_kernel_ms(lbg_ctx* ctx, double* ms, int* launches);
/* Device time (ms) of the dominant kernel alone in the last lbg_n0_ho02 call (the integrand kernel, n0_rows3_kernel by
 * default): CUDA events on the context's stream right around that one launch.  Numerator of bench.py's roofline. */
int         lbg_last_main_kernel_ms(lbg_ctx* ctx, double* ms);
/* Device time (CUDA events on the context stream) and kernel launches summed over every call since the last reset. */
int         lbg_kernel_totals(lbg_ctx* ctx, double* ms, long long* launches, int reset);
/* Dependent-DFMA micro-benchmark: measured FP64 FMA throughput of this GPU in TFLOP/s (FMA = 2 flop).
 * This is the denominator of the N0 roofline. */
int         lbg_fp64_peak(lbg_ctx* ctx, double* tflops);

/* ---- CMB spectra tables: replaces the dict of cubic interp1d objects built by prep_camb.py:86-89 ------- */
/* C_l on integer knots l = 0..n_ell-1, raw and dimensionless, order TT,EE,BB,TE (prep_camb.py:84).
 * The library builds the not-a-knot cubic spline of scipy `interp1d(kind='cubic')` itself (value 0 outside
 * the table, bounds_error=False / fill_value=0.0). */
int lbg_upload_cl_tables(lbg_ctx* ctx, int n_ell, const double* lensed /*[4][n_ell]*/,
                         const double* unlensed /*[4][n_ell]*/);
/* Evaluate the uploaded splines (prep_camb.py:88-89 callables): which = 0 lensed / 1 unlensed, spec = 0..3. */
int lbg_eval_cl(lbg_ctx* ctx, int which, int spec, int n, const double* x, double* out);

/* One-shot scipy `interp1d(arange(n_knots), y, kind='cubic', bounds_error=False, fill_value=0.0)(x)`. */
int lbg_eval_spline(lbg_ctx* ctx, int n_knots, const double* y, int n, const double* x, double* out);
/* Host-only (no GPU): the n_knots+2 uniform cubic B-spline coefficients c[-1..n_knots] of that spline. */
int lbg_cubic_spline_coeffs(int n_knots, const double* y, double* coef_out /*[n_knots+2]*/);

/* ---- N0, root flavour: replaces cmb/Nkk.py:54-172 (Nkk), :13-26 (prep_lmaxmask), prep_camb.py:128-137
 *      (Clxy) and cmb/DetectorNoise.py:8-32 ------------------------------------------------------------------
 * Quadrature nodes/weights are explicit ("kernels take weights, never a rule"):
 *   l1[n_l1], w_l1[n_l1]         the |u| grid (reference: arange(1, 5000.5, 0.5)) and its Simpson weights
 *   cos_phi/sin_phi/w_phi[n_phi]  the phi grid (reference: arange(0, 2pi+dphi, dphi)) and its weights
 * integral(L) = sum_ij w_phi[j] w_l1[i] l1[i] * sum_terms f F mask ;  N0 = (L(L+1)/2)^2 (2pi)^2 / integral.
 * lmax_T / lmax_P / lmin_v: the 3000 / 5000 / 2 cuts of prep_lmaxmask (0 < lmin_v < lmax_T <= lmax_P).
 * fold_phi: 0 = evaluate every phi node; 1 = evaluate one node of each mirror pair (phi, 2pi-phi) (the grid must be
 * symmetric) with the range-walking kernel; 2 = mirror pairs with the generic per-node-mask kernel (a second implementation of the
 * same sums, for cross-checks).  lbg_n0_last_path() tells which integrand kernel the last call ran: 5 = range-walking
 * (n0_rows5_kernel; needs fold_phi = 1, cos phi_j strictly decreasing over the first half of the grid, tables that cover every reachable
 * |v| <= min(lmax_P, max L + max l1) and a beam narrow enough for the folded-noise fit, about 7.5 arcmin at lmax_P = 5000),
 * 3 = generic (n0_rows3_kernel; anything else).  Inputs neither kernel can serve (more than 508 phi nodes after folding, |v| beyond
 * 16380) return LBG_EINVAL with a message; there is no other kernel and no environment switch.
 * *nodes_executed (may be NULL) receives the number of (phi, l1, L) nodes of the rows the main kernel was launched for.
 * Rows below the first non-zero C_l (l1 < 2 for CAMB tables, where C_0 = C_1 = 0): there 1/C^tot is pure detector noise and the
 * reference's value inherits the last-bit rounding of scipy's spline routines (DESIGN.md "parity notes").  The library evaluates the
 * u-side splines of every row with l1 < lbg_n0_set_exact_below() (default 2.0) on the host with scipy's own operation order
 * (make_interp_spline's LAPACK-order solve + de Boor evaluation, bit-identical: lbg_scipy_cubic_eval_host), so a C caller reaches the
 * reference to 1e-10 without scipy.  exact_rows[n_exact] / exact_u[n_exact][n_phi][8] (optional, n_exact may be 0): rows for which
 * the caller supplies its own values at |u|_j = sqrt((l cos_j)^2 + (l sin_j)^2) instead, order unlensed TT,TE,EE,BB then lensed
 * TT,TE,EE,BB.  Rows on the lmax cuts are always treated exactly.
 * out_terms (may be NULL) receives the per-estimator integrals, [5][n_L] in bit order TT,TE,EE,EB,BB. */
int lbg_n0_ho02(lbg_ctx* ctx, int n_L, const double* L, unsigned terms_mask,
                double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                int n_l1, const double* l1, const double* w_l1,
                int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                int n_exact, const int* exact_rows, const double* exact_u,
                double* out_nkk /*[n_L]*/, double* out_terms /*[5][n_L] or NULL*/,
                long long* nodes_executed /*or NULL*/);
/* The same call in two halves: _begin validates, stages the inputs in pinned memory and enqueues every kernel and the result copy on
 * the context's stream, then returns; _end waits and hands the results over.  Between the two the host (and other contexts of the
 * same GPU, lbg_ctx_create_lane) may do anything except call N0 / table uploads on THIS context (LBG_ESTATE).  A forecast starts
 * its 25 ms N0 curve first and builds the Limber spectra meanwhile -- one thread, no polling.  *nodes_unmasked (may be NULL): the
 * number of integrand nodes that passed a mask, counted by the kernels (the nodes the reference's masks leave non-zero). */
int lbg_n0_ho02_begin(lbg_ctx* ctx, int n_L, const double* L, unsigned terms_mask,
                      double thetab_arcmin, double deltaT_uK_arcmin, double Tcmb_K, int iterative,
                      int n_l1, const double* l1, const double* w_l1,
                      int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                      double lmax_T, double lmax_P, double lmin_v, int fold_phi,
                      int n_exact, const int* exact_rows, const double* exact_u);
int lbg_n0_ho02_end(lbg_ctx* ctx, int n_L, double* out_nkk /*[n_L]*/, double* out_terms /*[5][n_L] or NULL*/,
                    long long* nodes_executed /*or NULL*/, long long* nodes_unmasked /*or NULL*/);
/* Several noise levels of ONE beam in one pass (survey-design sweeps, BASELINE configs[4]; the reference loops Nkk over experiments,
 * snr.py:108-124): estimators TT,TE,EE,EB, folded phi grid.  For levels that share the beam C^tot = S + n_level G(l), so the lensed
 * splines and the noise shape are read once per node for all levels (n0_rows5m_kernel, groups of <= 3 levels): about half the
 * device time per level of lbg_n0_ho02.  deltaT_uK_arcmin[n_lev], iterative[n_lev] (may be NULL: all 0).  out_nkk[n_lev][n_L],
 * out_terms[n_lev][5][n_L] (may be NULL).  A level's result does not depend on which other levels travel with it; it agrees with
 * lbg_n0_ho02 to ~1e-13 (the noise enters after the Horner sums instead of inside the coefficients).  Inputs the range-walking kernel
 * cannot serve (see lbg_n0_ho02) return LBG_EINVAL: call lbg_n0_ho02 per level.  lbg_n0_last_path() reports 6. */
int lbg_n0_ho02_levels(lbg_ctx* ctx, int n_L, const double* L, double thetab_arcmin, int n_lev, const double* deltaT_uK_arcmin,
                       const int* iterative, double Tcmb_K, int n_l1, const double* l1, const double* w_l1,
                       int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                       double lmax_T, double lmax_P, double lmin_v, double* out_nkk, double* out_terms);
int lbg_n0_ho02_levels_begin(lbg_ctx* ctx, int n_L, const double* L, double thetab_arcmin, int n_lev, const double* deltaT_uK_arcmin,
                             const int* iterative, double Tcmb_K, int n_l1, const double* l1, const double* w_l1,
                             int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                             double lmax_T, double lmax_P, double lmin_v);
int lbg_n0_ho02_levels_end(lbg_ctx* ctx, int n_lev, int n_L, double* out_nkk, double* out_terms);
int lbg_n0_set_exact_below(lbg_ctx* ctx, double l_below);   /* 0 disables the scipy-order rows */
int lbg_n0_last_path(lbg_ctx* ctx, int* path);
/* Host-only (no GPU): scipy `interp1d(arange(n_knots), y, kind='cubic', bounds_error=False, fill_value=0.0)(x)`, bit for bit
 * (prep_camb.py:88-89 as scipy evaluates it: LAPACK-order not-a-knot solve, de Boor recurrence). */
int lbg_scipy_cubic_eval_host(int n_knots, const double* y, int n, const double* x, double* out);

/* ---- N0, modi flavour: replaces modi/nkk.py:35-92 (N_fisher: mode = LBG_TT, LBG_EE, LBG_EB) and :94-146 (N_fisherTE:
 *      mode = LBG_TE).  out_int[n_L] is the double integral `int2` those functions return;
 *      int2(L) = sum_i w_l[i] l[i] sum_j w_phi[j] frac(L, l_i, phi_j), frac = 0 where |l - L| lies outside [min l, max l].
 * Four spline tables on unit-spaced knots x0[t] + 0..tab_n[t]-1, values concatenated in tab_y; they are the
 * InterpolatedUnivariateSpline(k=3) arguments of the reference, in its argument order:
 *   TT/EE/EB: icl1, icl2 (unlensed), icln1, icln2 (lensed + detector noise);  TE: iclte, iclnt, iclne, iclnte.
 * The splines extrapolate beyond their knots (FITPACK ext=0), as the reference's do. */
int lbg_n0_modi(lbg_ctx* ctx, int mode, int n_L, const double* L, int n_l, const double* l, const double* w_l,
                int n_phi, const double* cos_phi, const double* sin_phi, const double* w_phi,
                const int* tab_n /*[4]*/, const double* tab_x0 /*[4]*/, const double* tab_y /*concatenated*/,
                double* out_int /*[n_L]*/);
/* scipy InterpolatedUnivariateSpline(x0 + arange(n_knots), y, k=3)(x) (modi/nkk.py:205-213 callables). */
int lbg_eval_spline_ext(lbg_ctx* ctx, int n_knots, double x0, const double* y, int n, const double* x, double* out);

/* ---- matter power: linear table x growth^2 -> HALOFIT; replaces modi/cosmology.py:158-182 (pkalin/pkanlin) and the
 *      array work of modi/halofit.py (sigma^2(R) integrals :36-44, Takahashi/Smith formula :139-212) ----------------- */
/* k [h/Mpc] increasing, P_lin(k, z=0) [(Mpc/h)^3], w_lnk = scipy-1.2.1 simps weights of the ln k grid (halofit.py:42). */
int lbg_upload_pk_linear(lbg_ctx* ctx, int n_k, const double* k, const double* plin_z0, const double* w_lnk);
/* sigma_0^2(R) = simps(Delta_0^2(k) exp(-(kR)^2), ln k) at z = 0 (multiply by D(z)^2 for other redshifts). */
int lbg_halofit_sigma2(lbg_ctx* ctx, int n_R, const double* R, double* out_sigma2);
/* Build the device-resident P(k, z_r) table, one row per entry of halo_z[n_rows][8] =
 * {D^2, k_nl, n_eff, curvature, Omega_m(z), Omega_L(z), use_halofit (0 -> linear row, modi/cosmology.py:171), unused}. */
int lbg_pk_build_table(lbg_ctx* ctx, int n_rows, const double* halo_z, double fnu, double om0, double w, int takahashi,
                       double* out_pnl /*[n_rows][n_k] or NULL: table stays on the device*/);
/* modi/halofit.py:70-98 (`_get_spec`'s two quintic InterpolatedUnivariateSpline fits, the inverse evaluation at ln sigma^2 = 0 and the
 * two derivatives) on the host in FITPACK's own operation order -- coefficients, values and derivatives are bit-identical to scipy's
 * (tests/test_fitpack_cpu.py), which matters because the curvature is the second derivative of an interpolating spline on a 0.0092
 * grid (rounding of the fit x 1e4).  lnr increasing, lnsig = ln sigma^2 on it (decreasing).  out = (k_nl = 1 / R_nl [the reference's `rknl`], n_eff, curvature); host-only.
 * *status: 0 ok, 1 = the derivative point ln(1 / rknl) lies outside the fitted range (scipy raises there: the reference's `except` branch, :79-96).
 * rknl_in > 0: skip the inverse fit and differentiate at ln(1 / rknl_in) (the refit inside that branch). */
int lbg_halofit_spec_host(int n, const double* lnr, const double* lnsig, double rknl_in, double* out /*[3]*/, int* status);
/* Number of k nodes of the linear table installed on the context (0: none). */
int lbg_pk_table_nk(lbg_ctx* ctx, int* n_k);
/* The same for whole z grids on the device, and the P(k,z) tables built from it (modi/cosmology.py:171-179 per node) without a host
 * round trip, for several z grids in one call: table i has n_rows[i] rows and is written to table slot slots[i] (lbg_pk_select_slot
 * numbering, -1 = the current slot); zin = the rows of all tables one after another, zin[row][4] = {D(z)^2, Omega_m(z), Omega_L(z),
 * use_halofit}; lnr[n_R] / R[n_R] = the reference's fitting grid np.linspace(ln 0.1, ln 10, 500) and its exp (modi/halofit.py:36-37).
 * sigma_0^2(R) is summed once per linear table in the reference's own order (scipy-1.2.1 simps panels + numpy's pairwise np.sum),
 * R_nl(z) comes from one inverse fit at z = 0 evaluated at -ln D^2 (the interpolant commutes with the shift; 1e-14), and (n_eff, C)
 * from a per-redshift fit of ln(D^2 sigma_0^2) in FITPACK order, one device thread per redshift on a factorisation shared by all.
 * out_spec (may be NULL): [rows][4] = {k_nl = 1 / R_nl, n_eff, curvature, flag}; flag 1 = R_nl outside the tabulated decade: that row
 * needs the reference's exception path (lbg_halofit_spec_host on a re-sampled grid + lbg_pk_build_table).  *n_flagged = number of
 * such rows.  build = 0: parameters only; 1: the tables are built when no row is flagged (row r = P(k, z_r)) and waited for;
 * 2: the build is enqueued and the call returns at once (*n_flagged = -1, no host outputs): the caller does its host work (evaluating
 * p(z), b(z)) meanwhile, later calls on the context are ordered behind the build, and lbg_pk_build_wait() reports the flagged count
 * (> 0: flagged rows were built as linear rows; rebuild through the explicit route before using results).  out_pnl: n_tables == 1. */
int lbg_pk_build_tables_hf(lbg_ctx* ctx, int n_tables, const int* n_rows, const int* slots, const double* zin,
                           int n_R, const double* lnr, const double* R, double fnu, double om0, double w, int takahashi, int build,
                           double* out_spec, int* n_flagged, double* out_pnl /*[n_rows][n_k] or NULL*/);
int lbg_pk_build_wait(lbg_ctx* ctx, int* n_flagged);
/* Enqueues (without waiting) the sigma_0^2(R) sums of the linear table just installed; lbg_pk_build_tables_hf picks them up.  Optional:
 * lets the caller overlap host work with the sums. */
int lbg_pk_prepare_hf(lbg_ctx* ctx, int n_R, const double* lnr, const double* R);
/* lbg_upload_pk_linear with the caller's own ln k values (numpy's np.log(k); NULL = this libm's): the sigma^2(R) sums of
 * lbg_pk_build_tables_hf then use exactly the reference's panel widths. */
int lbg_upload_pk_linear2(lbg_ctx* ctx, int n_k, const double* k, const double* plin_z0, const double* w_lnk, const double* lnk);
/* The context holds up to 4 P(k,z) tables; one is current (the one lbg_pk_build_table fills and lbg_limber / lbg_pk_eval read).
 * Makes table `slot` (0..3) current and returns its row count in *out_rows (0: nothing built in it yet).  Pointer swaps only.
 * Every table is dropped when a new linear table or k grid is installed (lbg_upload_pk_linear, lbg_eh_linear_table,
 * lbg_pk_upload_rows).  The reference has no counterpart: its camb interpolator answers any z (prep_camb.py:55-68); here a
 * forecast that alternates between the C_gg, C_kg and C_kk z grids keeps one table per grid instead of rebuilding. */
int lbg_pk_select_slot(lbg_ctx* ctx, int slot, int* out_rows);
/* Caller-supplied rows P_r(k) on their own k grid (the `kp = (k, P)` tables modi/lensing.py:127-150 interpolates): replaces
 * the device table built by lbg_pk_build_table (call lbg_upload_pk_linear again before rebuilding one). */
int lbg_pk_upload_rows(lbg_ctx* ctx, int n_k, const double* k, int n_rows, const double* rows /*[n_rows][n_k]*/);
/* np.interp(x, k, P_row) : oob = 0 -> 0 outside the k range (left=right=0, modi/lensing.py:175), 1 -> clamp (:161). */
int lbg_pk_eval(lbg_ctx* ctx, int row, int n, const double* x, int oob, double* out);

/* ---- Eisenstein & Hu (1999) MDM transfer function; replaces EisensteinHu/EisensteinAndHu.c:101-217 (TFmdm_set_cosm, arguments in
 *      its order: omega_matter, omega_baryon, omega_hdm, degen_hdm, omega_lambda, hubble [100 km/s/Mpc], redshift), :221-264
 *      (TFmdm_onek_mpc) and :268-280 (TFmdm_onek_hmpc: k in h/Mpc).  float64 and stateless where the reference is float32 with
 *      globals; same clamps (omega_baryon / omega_hdm <= 0 -> 1e-5, degen_hdm < 1 -> 1), theta_cmb = 2.7255 / 2.7 (:122).
 *      out_tf_cb = T_cb(k) (the reference's return value), out_tf_cbnu = its global tf_cbnu (may be NULL),
 *      out_growth_to_z0 = its global growth_to_z0 (may be NULL).  Illegal hubble / redshift (the reference exit(1)s) -> LBG_EINVAL. */
int lbg_eh_transfer(lbg_ctx* ctx, double omega_matter, double omega_baryon, double omega_hdm, int degen_hdm, double omega_lambda,
                    double hubble, double redshift, int n_k, const double* k_hmpc, double* out_tf_cb, double* out_tf_cbnu,
                    double* out_growth_to_z0);
/* Linear table P(k, z=0) = A k^ns T_cb(k)^2 with A fixed by sigma8 (top-hat R = 8 Mpc/h, trapezoid in ln k over the grid), built on
 * the device and left on the context exactly as lbg_upload_pk_linear leaves a host table (cosmology sweeps
 * need no host table).  w_lnk as in lbg_upload_pk_linear; out_plin (may be NULL) receives the table. */
int lbg_eh_linear_table(lbg_ctx* ctx, double omega_matter, double omega_baryon, double omega_hdm, int degen_hdm, double omega_lambda,
                        double hubble, double ns, double sigma8, int n_k, const double* k, const double* w_lnk, double* out_plin);

/* ---- linear growth tables; replaces the 1000 `romberg` calls of modi/cosmology.py:92-111 (_eval_Dplus) for n_cosmo cosmologies in
 *      one launch: out_y[c][i] = E(a_i) * romberg(kernel, loga[0], loga[i]) with kernel(ln a) = (a E(a))^-3 a and
 *      E(a) = a^-1.5 sqrt(M + (1 - M - L) a + L a^3) (modi/cosmology.py:119-125); the scipy-1.2.1 rule (tol = rtol = 1.48e-8,
 *      at most 2^10 panels).  D(a) / D(1) is then the reference's cubic spline through (a_i, y_i) (:111-117), left to the host.
 *      out_levels (may be NULL) receives the refinement level at which each node stopped. */
int lbg_growth_table(lbg_ctx* ctx, int n_cosmo, const double* omega_m, const double* omega_l, int n, const double* loga,
                     double* out_y, int* out_levels);

/* ---- Limber: C_l[b][l] = sum_z weight[b][z] * P_{prow[z]}(k = (l + 1/2)/chi[z]); replaces the z loops + quadrature of
 *      Cgg.py:53-68, Ckg.py:42-59, cmb/lensing.py:27-43 (Simpson in z) and modi/lensing.py:127-179 (trapezoid in chi).
 *      weight = kernel product x bias x unit factors x quadrature weight, so one kernel serves every flavour. */
int lbg_limber(lbg_ctx* ctx, int n_batch, int n_ell, const double* ell, int n_z, const double* chi,
               const double* weight /*[n_batch][n_z]*/, const int* prow /*[n_z]*/, int oob, double* out_cl /*[n_batch][n_ell]*/);

/* ---- the Limber stage of one forecast in one call: interloper.py:23-40 (`get_allCls`) without its N0 call, for n_s galaxy samples
 *      on one set of multipoles: C_gg (Cgg.py:53-68, zeff=False, auto spectrum), C_kg (Ckg.py:42-59, zeff=False) per sample and
 *      C_kk (cmb/lensing.py:21-52) once.  The host only evaluates the callables: pz_*[s][n] = p_s(z) and bz_*[s][n] = b_s(z) on the
 *      C_gg nodes (Cgg.py:18-19) and on the C_kg nodes (Ckg.py:19-20), norm[s] = trapz(p_s) on the dz = 0.001 grid of sliced_pz.py:6-9.
 *      On the device: the sliced_pz renormalisation (sliced_pz.py:10-26, dz_* = its 6-decimal rounded grid spacing), the prefactors
 *      (H/c) p^ p^ / chi^2 b b / h and W_kappa p^ / chi^2 b / h^2 times the Simpson weights w_*, and the sums over z against the
 *      P(k,z) tables parked in slot_gg / slot_kg / slot_kk (lbg_pk_select_slot + lbg_pk_build_table: one row per grid node, in order).
 *      hc_gg = H(z)/c on the C_gg nodes, wk_kg = W_kappa(z) on the C_kg nodes (cmb/lensing.py:14-19), weight_kk = the complete
 *      per-node weight of the C_kk sum (n_kk = 0: no C_kk).  One pinned staging block in, one out. */
int lbg_forecast_spectra(lbg_ctx* ctx, int n_ell, const double* ell, double zmin, double zmax, double h100,
                         int n_gg, const double* z_gg, const double* chi_gg, const double* hc_gg, const double* w_gg, double dz_gg, int slot_gg,
                         int n_kg, const double* z_kg, const double* chi_kg, const double* wk_kg, const double* w_kg, double dz_kg, int slot_kg,
                         int n_kk, const double* chi_kk, const double* weight_kk, int slot_kk,
                         int n_s, const double* pz_gg, const double* bz_gg, const double* pz_kg, const double* bz_kg, const double* norm,
                         double* out_cgg /*[n_s][n_ell]*/, double* out_ckg /*[n_s][n_ell]*/, double* out_ckk /*[n_ell] or NULL*/);

/* ---- S/N, covariance and Fisher accumulation over multipoles ---------------------------------------------------- */
/* snr.py:21-25 `snr`: sqrt( sum_{l <= lmax} nmodes (C_l / sqrt(Var_l))^2 ). */
int lbg_snr(lbg_ctx* ctx, int n_ell, const double* ell, const double* cl, const double* var, const double* nmodes,
            double lmax, double* out_snr);
/* The experiment x delensing x dilution sweep of snr.py:107-133 as ONE call: for experiment e (N_kk row e, fsky[e]), delensing curve r
 * (rho[r][l]; a row of zeros = no delensing) and shot noise ngg[d] (= Ngg of the diluted density, Cgg.py:85-95),
 *   out[e][r][d] = snr(L, C_kg, var, nmodes, lmax)   with   var = (((1 - rho^2) C_kk + N_kk)(C_gg + N_gg) + C_kg^2) / ((2L + 1) fsky)
 * (Ckg.py:61-87 incl. the rho branch :69-71, snr.py:21-25).  Bit-identical to var_Ckg + snr called per design.  The caller divides
 * by sqrt(fsky) where the reference driver does (snr.py:125). */
int lbg_snr_grid(lbg_ctx* ctx, int n_ell, const double* ell, const double* nmodes, const double* ckk, const double* cgg, const double* ckg,
                 int n_exp, const double* nkk /*[n_exp][n_ell]*/, const double* fsky /*[n_exp]*/, int n_rho, const double* rho /*[n_rho][n_ell]*/,
                 int n_dil, const double* ngg /*[n_dil]*/, double lmax, double* out /*[n_exp][n_rho][n_dil]*/);
/* lbg_snr for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * cudaMalloc).  No counterpart in the reference.  Same arguments, same kernel and same
 * result as lbg_snr, but the device staging of the four vectors and of the result lives in a fixed file-scope device array instead of
 * the context's grow-only scratch: the call performs no cudaMalloc / cudaFree / allocating library call.  Limit: n_ell <= 16384
 * (LBG_EINVAL beyond, with a message); the array (512 KiB + 256 B) is held by every GPU that loads the library until the process
 * exits.  All calls on a device share it and it keeps the previous call's contents (the result slot is cleared per call): calls that
 * overlap on one GPU would race, so the library serialises them with a process-wide lock. */
int lbg_snr_noalloc(lbg_ctx* ctx, int n_ell, const double* ell, const double* cl, const double* var, const double* nmodes,
                    double lmax, double* out_snr);
/* interloper.py:42-63 `get_ClsCov` for any set of fields: tot[l][a][b] = (C + N)^{ab}(l) (n_f x n_f, symmetric);
 * spectrum i is the field pair (spec_a[i], spec_b[i]);
 * cov[l][i][j] = (T^{ac} T^{bd} + T^{ad} T^{bc}) / (2 l + 1) / fsky  with (a,b) = pair i and (c,d) = pair j. */
int lbg_gaussian_cov(lbg_ctx* ctx, int n_ell, const double* ell, int n_f, const double* tot /*[n_ell][n_f][n_f]*/,
                     int n_spec, const int* spec_a, const int* spec_b, double fsky, double* out_cov /*[n_ell][n_spec][n_spec]*/);
/* interloper.py:223-253 (Fisher matrix and bias vector) and yu.py:99-106 (rho_L^2 = r^T rho^-1 r, n_par = 1):
 * per multipole Q_l[p][q] = dD_p^T Cov_l^-1 dD_q and b_l[p] = dD_p^T Cov_l^-1 dDelta_l (LU with partial pivoting on the
 * diagonally equilibrated matrix); out_F = sum_l mode_weight[l] Q_l, out_b = sum_l mode_weight[l] b_l (fixed-order sums).
 * n_spec <= 32, n_par <= 8.  dDelta / out_b / out_per_ell ([n_ell][n_par][n_par], the Q_l themselves) may be NULL.
 * A multi-GPU caller shards l, calls this per rank and all-reduces out_F / out_b (lbgcmb_b200/dist.py, NCCL). */
int lbg_fisher_accum(lbg_ctx* ctx, int n_ell, int n_spec, int n_par, const double* cov /*[n_ell][n_spec][n_spec]*/,
                     const double* dD /*[n_ell][n_par][n_spec]*/, const double* dDelta /*[n_ell][n_spec] or NULL*/,
                     const double* mode_weight /*[n_ell]*/, double* out_F /*[n_par][n_par]*/, double* out_b /*[n_par] or NULL*/,
                     double* out_per_ell /*or NULL*/);
/* The same Fisher sum for n_f correlated fields at once, in field form (no n_spec x n_spec covariance is formed):
 *   out_F[p][q] = sum_l mode_weight[l] * 1/2 Tr[ C_l^-1 d_pC_l  C_l^-1 d_qC_l ],  C_l[a][b] = (C + N)^{ab}(l) symmetric, n_f x n_f,
 * which equals interloper.py:223-253 (`dD^T Cov^-1 dD` with the interloper.py:42-63 covariance) when the data vector holds every auto
 * and cross spectrum of the fields and mode_weight[l] = nmodes[l] (2 l + 1) fsky.  Spectrum s is the field pair (spec_a[s], spec_b[s]);
 * pairs that are not listed are zero in C_l and in its derivatives (disjoint redshift bins); every field needs its auto spectrum.
 * tot[n_spec][n_ell] = C + N, dC[n_par][n_spec][n_ell] = parameter derivatives.  n_f <= 160, n_par <= 8 (BASELINE configs[3]:
 * kappa + 6 tracers x 20 bins = 121 fields, 241 spectra).  out_per_ell ([n_ell][n_par][n_par], unweighted) may be NULL. */
int lbg_field_fisher(lbg_ctx* ctx, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, const double* tot,
                     int n_par, const double* dC, const double* mode_weight, double* out_F, double* out_per_ell /*or NULL*/);
/* lbg_field_fisher with amplitude-type parameters described instead of tabulated: parameters 0..n_dense-1 have explicit rows
 * dC[n_dense][n_spec][n_ell]; parameter n_dense + a is an amplitude A_a with d C_s / d A_a = amp[a][s] * sig[s][l] (sig[n_spec][n_ell] = the
 * signal spectra C without noise; e.g. a bias amplitude: amp = 1 on its kappa-g spectra, 2 on its auto spectra, 0 elsewhere -- the
 * analytic dp_DL('b1') of interloper.py:65-74).  Their rows are formed on the device: a tomographic forecast with 2 cosmological + 6 bias
 * parameters uploads 38 MB instead of 86 MB and the host never builds the 57 MB of copied spectra.  n_dense + n_amp <= 8. */
int lbg_field_fisher_amp(lbg_ctx* ctx, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, const double* tot,
                         int n_dense, const double* dC, int n_amp, const double* amp, const double* sig,
                         const double* mode_weight, double* out_F, double* out_per_ell /*or NULL*/);
/* Spectra kept on the device between calls, for forecasts whose Fisher inputs are differences of Limber sums (the host loop of
 * interloper.py:223-253 taken over configs[3]'s 241 spectra x 5 cosmologies): a context holds up to 20 spectra sets, each a
 * [set_rows][n_ell] matrix.
 *   lbg_limber_keep       lbg_limber whose batch row b stays on the device as row dest_row[b] of set `set` (dest_row[b] < 0: dropped); the
 *                         first call that names a set with a new shape zeroes it, later calls fill further rows.
 *   lbg_spectra_set_fill  a set filled without a Limber sum (noise): row r = row_const[r] at every multipole, then rows curve_row[i]
 *                         replaced by curves[i][n_ell] (n_curves may be 0).
 *   lbg_spectra_set_get   rows [row0, row0 + n_rows) of a set copied to the host.
 *   lbg_field_fisher_sets lbg_field_fisher_amp with tot = fid + noise (set_noise < 0: none), explicit derivative rows
 *                         dC_j = (plus_j - minus_j) / two_h[j] (j < n_diff; the roundings of numpy's subtract-then-divide) and sig = fid, all
 *                         formed on the device from the named sets (each n_spec rows x n_ell).  n_diff + n_amp <= 8.
 * No counterpart in the reference (its spectra live in numpy arrays); the results equal the host-array entry points bit for bit. */
int lbg_limber_keep(lbg_ctx* ctx, int n_batch, int n_ell, const double* ell, int n_z, const double* chi, const double* weight,
                    const int* prow, int oob, int set, int set_rows, const int* dest_row /*[n_batch]*/);
int lbg_spectra_set_fill(lbg_ctx* ctx, int set, int set_rows, int n_ell, const double* row_const /*[set_rows]*/, int n_curves,
                         const int* curve_row, const double* curves /*[n_curves][n_ell]*/);
int lbg_spectra_set_get(lbg_ctx* ctx, int set, int row0, int n_rows, int n_ell, double* out /*[n_rows][n_ell]*/);
int lbg_field_fisher_sets(lbg_ctx* ctx, int n_ell, int n_f, int n_spec, const int* spec_a, const int* spec_b, int set_fid, int set_noise,
                          int n_diff, const int* set_plus, const int* set_minus, const double* two_h, int n_amp, const double* amp,
                          const double* mode_weight, double* out_F, double* out_per_ell /*or NULL*/);
/* ---- the exchange step of a forecast whose multipoles are sharded over the GPUs of one box, over NVLink peer memory.
 * One process per GPU.  lbg_xchg_create allocates this rank's window (room for blocks of up to cap_doubles doubles from every rank,
 * double-buffered) and returns its 64-byte CUDA IPC handle; the caller passes the handles of all ranks around (any transport:
 * torch.distributed in lbgcmb_b200/dist.py) and hands them to lbg_xchg_open.  lbg_xchg_allgather is then a collective: every rank
 * calls it the same number of times with the same n; block[n] of rank r arrives as out[r][n] on every rank -- one kernel per rank
 * that stores the block into every peer's window and waits for the peers' blocks (release / acquire flags), no library collective.
 * A rank that does not arrive within ~4 s of GPU clocks ends the call with LBG_ECUDA instead of hanging.  *out_ms: device time of
 * the kernel (includes the wait for the slowest peer).  No counterpart in the reference (single process). */
int lbg_xchg_create(lbg_ctx* ctx, int rank, int world, long long cap_doubles, unsigned char* out_handle /*[64]*/);
int lbg_xchg_open(lbg_ctx* ctx, const unsigned char* handles /*[world][64], rank order*/);
int lbg_xchg_allgather(lbg_ctx* ctx, long long n, const double* block /*[n]*/, double* out /*[world][n]*/, double* out_ms /*or NULL*/);
/* The exchange step of a sharded forecast in one call (what lbgcmb_b200/dist.py `exchange` makes): sums[n_sums] are added over
 * the ranks, in rank order (the all-reduce of the Fisher / S/N^2 / bias sums of interloper.py:223-253, snr.py:21-25 taken over
 * multipole shards), and curves[n_curves][n_mine] -- this rank's cyclic shard, units rank, rank + world, ... of n -- come back whole
 * as out_curves[n_curves][n]. */
int lbg_xchg_sums_curves(lbg_ctx* ctx, int n_sums, const double* sums, int n_curves, long long n, const double* curves,
                         double* out_sums, double* out_curves, double* out_ms /*or NULL*/);
int lbg_xchg_destroy(lbg_ctx* ctx);
/* mqw/py/mqw.py:63-104 `Fisher` (clustering-redshift N_p(z_i) Fisher): bins[n_bin][8] = {bp, bs, Np, Ns, wp, ws, wps, -}
 * per redshift shell (mqw.py:24-41); cgg[n_ell] shared by all shells (cgg_per_bin = 0, as mqw.py:48) or
 * cgg[n_bin][n_ell] (cgg_per_bin = 1, as cadence/cadence.py:97); mode_weight[l] = 2l+1 (0 beyond an l cut).
 * out_F[n_bin][n_bin]: mqw.py:97-104; out_Fsl[n_bin]: the Schur-Limber diagonal F_ii of mqw.py:83; out_S[n_ell]: S(l). */
int lbg_mqw_fisher(lbg_ctx* ctx, int n_ell, int n_bin, const double* cgg, int cgg_per_bin, const double* bins /*[n_bin][8]*/,
                   const double* mode_weight /*[n_ell]*/, double fsky, double* out_F, double* out_Fsl /*or NULL*/,
                   double* out_S /*or NULL*/);

/* cadence/cadence.py:64-191 `Fisher` (the LSST-lens flavour of the same forecast): per-shell cgg[n_bin][n_ell], the same bins
 * layout, mode_weight[l] = 2l+1 inside the reference's LCUT (cadence.py:73-74: the cut of the LAST shell) and 0 beyond.
 * out_F reproduces cadence.py:155-165 as written, including its running `Interim` sum:
 *   F_ij = [i == j] fsky B_i + B_i + sum_{j' <= j} T_ij',  B_i = sum_l w S A0i_j^2 / Aii / A00,  T_ij = the cross term of mqw.py:100. */
int lbg_cadence_fisher(lbg_ctx* ctx, int n_ell, int n_bin, const double* cgg /*[n_bin][n_ell]*/, const double* bins /*[n_bin][8]*/,
                       const double* mode_weight /*[n_ell]*/, double fsky, double* out_F, double* out_Fsl /*or NULL*/,
                       double* out_S /*or NULL*/);

/* ---- survey-design grid (BASELINE configs[4]): for every (noise level n, galaxy configuration g) the cumulative S/N of
 *      C_kg (snr.py:118-124 with the Ckg.py:83-87 variance) and the (sigma8, b1) Fisher matrix + interloper bias vector of
 *      D = (C_kk, C_kg, C_gg) (interloper.py:200-253) -- the same per-multipole arithmetic as lbg_gaussian_cov +
 *      lbg_fisher_accum + lbg_snr, one CTA per design.
 *      cgg/ckg[n_cfg][n_ell]; xcgg/xckg: interloper-contaminated spectra (both NULL: no bias vector);
 *      cfg[n_cfg][8] = {N_gg, contaminated N_gg, fiducial b1, L cut of the Fisher sums (L < cut), L cut of the S/N (L <= cut), -,-,-};
 *      nkk[n_noise][n_ell] = N0 per noise level;  out[n_noise][n_cfg][8] = {S/N, F00, F01, F10, F11, b0, b1, 0}
 *      (S/N is not divided by sqrt(fsky); parameters in the order sigma8, b1). */
int lbg_forecast_grid(lbg_ctx* ctx, int n_ell, const double* ell, const double* nmodes, const double* ckk,
                      int n_cfg, const double* cgg, const double* ckg, const double* xcgg, const double* xckg,
                      const double* cfg, int n_noise, const double* nkk, double fsky, double fid_sig8, double* out);

#ifdef __cplusplus
}
#endif
#endif /* LBGCMB_H */
