"""Device-side covariance / Fisher primitives shared by the mirrors of interloper.py, mqw.py and yu.py
(csrc/fisher.cu through lbg_gaussian_cov, lbg_fisher_accum, lbg_mqw_fisher).  No CPU path."""
import numpy as np

from lbgcmb_b200 import _ffi


def gaussian_cov(Llls, tot, spectra, fsky, device=None):
    """cov[l, i, j] of interloper.py:42-63 for field pairs `spectra` = [(a, b), ...] and tot[l, a, b] = (C + N)^{ab}."""
    ctx = _ffi.context(device)
    L = _ffi.f64(Llls)
    tot = _ffi.f64(tot)
    if tot.ndim != 3 or tot.shape[0] != L.size or tot.shape[1] != tot.shape[2]:
        raise ValueError('tot must be [n_ell, n_f, n_f]')
    sa = np.ascontiguousarray([s[0] for s in spectra], dtype=np.int32)
    sb = np.ascontiguousarray([s[1] for s in spectra], dtype=np.int32)
    out = np.empty((L.size, sa.size, sa.size))
    _ffi.check(_ffi.lib().lbg_gaussian_cov(ctx.handle, L.size, _ffi._dp(L), tot.shape[1], _ffi._dp(tot), sa.size,
                                           sa.ctypes.data_as(_ffi.c_int_p), sb.ctypes.data_as(_ffi.c_int_p), float(fsky), _ffi._dp(out)))
    return out


def fisher_accum(cov, dD, mode_weight, dDelta=None, per_ell=False, device=None):
    """F = sum_l w_l dD_l Cov_l^-1 dD_l^T  (and b = sum_l w_l dD_l Cov_l^-1 dDelta_l); interloper.py:223-253.

    cov [n_ell, n_spec, n_spec]; dD [n_ell, n_par, n_spec]; dDelta [n_ell, n_spec] or None.
    Returns F, b (None without dDelta) and, with per_ell, the per-multipole Q_l [n_ell, n_par, n_par]."""
    ctx = _ffi.context(device)
    cov, dD, w = _ffi.f64(cov), _ffi.f64(dD), _ffi.f64(mode_weight)
    n_ell, n_par, n_spec = dD.shape
    if cov.shape != (n_ell, n_spec, n_spec) or w.shape != (n_ell,):
        raise ValueError('cov / dD / mode_weight shapes disagree')
    dl = None
    if dDelta is not None:
        dl = _ffi.f64(dDelta)
        if dl.shape != (n_ell, n_spec):
            raise ValueError('dDelta must be [n_ell, n_spec]')
    F = np.empty((n_par, n_par))
    b = np.empty(n_par) if dl is not None else None
    Q = np.empty((n_ell, n_par, n_par)) if per_ell else None
    _ffi.check(_ffi.lib().lbg_fisher_accum(ctx.handle, n_ell, n_spec, n_par, _ffi._dp(cov), _ffi._dp(dD), _ffi._dp(dl), _ffi._dp(w),
                                           _ffi._dp(F), _ffi._dp(b), _ffi._dp(Q)))
    return (F, b, Q) if per_ell else (F, b)


def field_fisher(n_f, spectra, tot, dC, mode_weight, per_ell=False, device=None):
    """F[p, q] = sum_l w_l 1/2 Tr[C_l^-1 d_pC_l C_l^-1 d_qC_l] for n_f correlated fields (lbg_field_fisher): the interloper.py:223-253
    Fisher sum over ALL auto and cross spectra of the fields (w_l = nmodes (2l+1) fsky), without the n_spec x n_spec covariance.

    spectra = [(a, b), ...] field pairs (unlisted pairs are zero); tot [n_spec, n_ell] = C + N; dC [n_par, n_spec, n_ell]."""
    ctx = _ffi.context(device)
    tot, dC, w = _ffi.f64(tot), _ffi.f64(dC), _ffi.f64(mode_weight)
    n_spec, n_ell = tot.shape
    if dC.ndim != 3 or dC.shape[1:] != (n_spec, n_ell) or w.shape != (n_ell,) or len(spectra) != n_spec:
        raise ValueError('spectra / tot / dC / mode_weight shapes disagree')
    n_par = dC.shape[0]
    sa = np.ascontiguousarray([s[0] for s in spectra], dtype=np.int32)
    sb = np.ascontiguousarray([s[1] for s in spectra], dtype=np.int32)
    F = np.empty((n_par, n_par))
    Q = np.empty((n_ell, n_par, n_par)) if per_ell else None
    _ffi.check(_ffi.lib().lbg_field_fisher(ctx.handle, n_ell, int(n_f), n_spec, sa.ctypes.data_as(_ffi.c_int_p),
                                           sb.ctypes.data_as(_ffi.c_int_p), _ffi._dp(tot), n_par, _ffi._dp(dC), _ffi._dp(w), _ffi._dp(F),
                                           _ffi._dp(Q) if per_ell else None))
    if not np.all(np.isfinite(F)):                     # a zero pivot: the reference's np.linalg.inv raises on a singular covariance
        raise np.linalg.LinAlgError('lbg_field_fisher: C_l is singular (zero pivot) or holds non-finite entries at some multipole')
    return (F, Q) if per_ell else F


def field_fisher_amp(n_f, spectra, tot, dC, amp, sig, mode_weight, device=None):
    """field_fisher with amplitude parameters described, not tabulated (lbg_field_fisher_amp): dC [n_dense, n_spec, n_ell] for the
    parameters with explicit derivatives, then one parameter per row of amp [n_amp, n_spec] with d C_s / d A = amp[a, s] * sig[s]
    (sig [n_spec, n_ell]: the signal spectra).  Parameter order of F: dense ones first."""
    ctx = _ffi.context(device)
    tot, sig, amp, w = _ffi.f64(tot), _ffi.f64(sig), _ffi.f64(np.atleast_2d(amp)), _ffi.f64(mode_weight)
    n_spec, n_ell = tot.shape
    dC = _ffi.f64(dC) if dC is not None and np.size(dC) else np.zeros((0, n_spec, n_ell))
    if dC.ndim != 3 or dC.shape[1:] != (n_spec, n_ell) or sig.shape != tot.shape or amp.shape[1] != n_spec or w.shape != (n_ell,) \
            or len(spectra) != n_spec:
        raise ValueError('spectra / tot / dC / amp / sig / mode_weight shapes disagree')
    n_par = dC.shape[0] + amp.shape[0]
    sa = np.ascontiguousarray([s_[0] for s_ in spectra], dtype=np.int32)
    sb = np.ascontiguousarray([s_[1] for s_ in spectra], dtype=np.int32)
    F = np.empty((n_par, n_par))
    _ffi.check(_ffi.lib().lbg_field_fisher_amp(ctx.handle, n_ell, int(n_f), n_spec, sa.ctypes.data_as(_ffi.c_int_p),
                                               sb.ctypes.data_as(_ffi.c_int_p), _ffi._dp(tot), dC.shape[0], _ffi._dp(dC), amp.shape[0],
                                               _ffi._dp(amp), _ffi._dp(sig), _ffi._dp(w), _ffi._dp(F), None))
    if not np.all(np.isfinite(F)):
        raise np.linalg.LinAlgError('lbg_field_fisher: C_l is singular (zero pivot) or holds non-finite entries at some multipole')
    return F


def spectra_set_fill(set_index, row_const, n_ell, curves=None, device=None):
    """Spectra set `set_index` on the device <- rows constant over the multipoles (row_const [n_rows]), rows named by `curves`
    ({row: [n_ell] array}) then replaced (lbg_spectra_set_fill): the noise matrix of a forecast without building it on the host."""
    ctx = _ffi.context(device)
    rc = _ffi.f64(row_const)
    curves = curves or {}
    rows = np.ascontiguousarray(sorted(curves), dtype=np.int32)
    cv = _ffi.f64(np.array([curves[r] for r in rows], dtype=np.float64).reshape(rows.size, int(n_ell)))
    _ffi.check(_ffi.lib().lbg_spectra_set_fill(ctx.handle, int(set_index), rc.size, int(n_ell), _ffi._dp(rc), rows.size,
                                               rows.ctypes.data_as(_ffi.c_int_p), _ffi._dp(cv)))


def spectra_set_get(set_index, row0, n_rows, n_ell, device=None):
    """Rows [row0, row0 + n_rows) of a kept spectra set (lbg_spectra_set_get)."""
    ctx = _ffi.context(device)
    out = np.empty((int(n_rows), int(n_ell)))
    _ffi.check(_ffi.lib().lbg_spectra_set_get(ctx.handle, int(set_index), int(row0), int(n_rows), int(n_ell), _ffi._dp(out)))
    return out


def field_fisher_sets(n_f, spectra, n_ell, set_fid, set_noise, diffs, amp, mode_weight, device=None):
    """field_fisher_amp whose spectra never left the device (lbg_field_fisher_sets): tot = set_fid + set_noise (None: no noise),
    explicit derivatives (set_plus - set_minus) / two_h for each (set_plus, set_minus, two_h) of `diffs`, sig = set_fid; the sets were
    filled by Cosmology.limber(..., keep=...) / spectra_set_fill."""
    ctx = _ffi.context(device)
    n_spec = len(spectra)
    amp = _ffi.f64(np.atleast_2d(amp)) if amp is not None and np.size(amp) else np.zeros((0, n_spec))
    w = _ffi.f64(mode_weight)
    if amp.shape[1] != n_spec or w.shape != (int(n_ell),):
        raise ValueError('spectra / amp / mode_weight shapes disagree')
    sa = np.ascontiguousarray([s_[0] for s_ in spectra], dtype=np.int32)
    sb = np.ascontiguousarray([s_[1] for s_ in spectra], dtype=np.int32)
    plus = np.ascontiguousarray([d_[0] for d_ in diffs], dtype=np.int32)
    minus = np.ascontiguousarray([d_[1] for d_ in diffs], dtype=np.int32)
    two_h = _ffi.f64([d_[2] for d_ in diffs])
    n_par = len(diffs) + amp.shape[0]
    F = np.empty((n_par, n_par))
    _ffi.check(_ffi.lib().lbg_field_fisher_sets(ctx.handle, int(n_ell), int(n_f), n_spec, sa.ctypes.data_as(_ffi.c_int_p),
                                                sb.ctypes.data_as(_ffi.c_int_p), int(set_fid), -1 if set_noise is None else int(set_noise),
                                                len(diffs), plus.ctypes.data_as(_ffi.c_int_p), minus.ctypes.data_as(_ffi.c_int_p),
                                                _ffi._dp(two_h), amp.shape[0], _ffi._dp(amp), _ffi._dp(w), _ffi._dp(F), None))
    if not np.all(np.isfinite(F)):
        raise np.linalg.LinAlgError('lbg_field_fisher: C_l is singular (zero pivot) or holds non-finite entries at some multipole')
    return F


def mqw_fisher(cgg, bins, mode_weight, fsky, device=None):
    """(F [n_bin, n_bin], F_ii Schur-Limber [n_bin], S(l)) of mqw/py/mqw.py:63-104; bins[n_bin, 7] = bp,bs,Np,Ns,wp,ws,wps."""
    ctx = _ffi.context(device)
    cgg, w = _ffi.f64(cgg), _ffi.f64(mode_weight)
    b = np.zeros((len(bins), 8))
    b[:, :7] = np.asarray(bins, dtype=np.float64)
    n_bin, n_ell = b.shape[0], w.size
    per_bin = cgg.ndim == 2
    if cgg.shape != ((n_bin, n_ell) if per_bin else (n_ell,)):
        raise ValueError('cgg must be [n_ell] or [n_bin, n_ell]')
    F, Fsl, S = np.empty((n_bin, n_bin)), np.empty(n_bin), np.empty(n_ell)
    _ffi.check(_ffi.lib().lbg_mqw_fisher(ctx.handle, n_ell, n_bin, _ffi._dp(cgg), int(per_bin), _ffi._dp(b), _ffi._dp(w), float(fsky),
                                         _ffi._dp(F), _ffi._dp(Fsl), _ffi._dp(S)))
    return F, Fsl, S
