"""ctypes binding of liblbgcmb.so (include/lbgcmb.h).  No CPU fallback: a missing library or GPU raises."""
import ctypes as C
import os
import threading

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get('LBGCMB_LIB') or os.path.join(_HERE, 'liblbgcmb.so')   # override: kernel experiments only

_lib = None
_lock = threading.Lock()
_ctxs = {}
# device index -> (token, digest) of the python object whose linear P(k) table currently sits on that context (None: nobody's).
# The token comes from `new_owner_token()` (never reused, unlike id()); the digest covers the table's bytes, so a table edited in
# place is uploaded again.
table_owner = {}
_owner_tokens = __import__('itertools').count(1)


def new_owner_token():
    return next(_owner_tokens)


def table_digest(*arrays):
    import hashlib
    h = hashlib.blake2b(digest_size=16)
    for a in arrays:
        a = np.ascontiguousarray(a, dtype=np.float64)
        h.update(str(a.shape).encode())
        h.update(a.tobytes())
    return h.digest()

c_double_p = C.POINTER(C.c_double)
c_ll_p = C.POINTER(C.c_longlong)
c_int_p = C.POINTER(C.c_int)


class LbgError(RuntimeError):
    pass


def _dp(a):
    return a.ctypes.data_as(c_double_p) if a is not None else None


def f64(a):
    """C-contiguous float64 copy/view of `a` (what every entry point expects)."""
    return np.ascontiguousarray(a, dtype=np.float64)


def lib():
    """Load liblbgcmb.so (built by lbgcmb_b200/csrc/build.sh or __graft_entry__.build())."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.isfile(LIB_PATH):
                raise LbgError('%s not found: run lbgcmb_b200/csrc/build.sh (there is no CPU fallback)' % LIB_PATH)
            L = C.CDLL(LIB_PATH)
            L.lbg_last_error.restype = C.c_char_p
            L.lbg_version.restype = C.c_char_p
            _declare(L)
            _lib = L
    return _lib


def _declare(L):
    vp, i, d, u = C.c_void_p, C.c_int, C.c_double, C.c_uint
    sig = {
        'lbg_ctx_create': [i, C.POINTER(vp)],
        'lbg_ctx_create_lane': [i, i, C.POINTER(vp)],
        'lbg_pk_select_slot': [vp, i, C.POINTER(C.c_int)],
        'lbg_growth_table': [vp, i, c_double_p, c_double_p, i, c_double_p, c_double_p, C.POINTER(C.c_int)],
        'lbg_ctx_destroy': [vp],
        'lbg_last_kernel_ms': [vp, c_double_p, c_int_p],
        'lbg_last_main_kernel_ms': [vp, c_double_p],
        'lbg_kernel_totals': [vp, c_double_p, c_ll_p, i],
        'lbg_fp64_peak': [vp, c_double_p],
        'lbg_upload_cl_tables': [vp, i, c_double_p, c_double_p],
        'lbg_eval_cl': [vp, i, i, i, c_double_p, c_double_p],
        'lbg_eval_spline': [vp, i, c_double_p, i, c_double_p, c_double_p],
        'lbg_cubic_spline_coeffs': [i, c_double_p, c_double_p],
        'lbg_upload_pk_linear': [vp, i, c_double_p, c_double_p, c_double_p],
        'lbg_halofit_sigma2': [vp, i, c_double_p, c_double_p],
        'lbg_pk_build_table': [vp, i, c_double_p, d, d, d, i, c_double_p],
        'lbg_halofit_spec_host': [i, c_double_p, c_double_p, d, c_double_p, c_int_p],
        'lbg_pk_table_nk': [vp, c_int_p],
        'lbg_pk_build_tables_hf': [vp, i, c_int_p, c_int_p, c_double_p, i, c_double_p, c_double_p, d, d, d, i, i, c_double_p, c_int_p, c_double_p],
        'lbg_pk_build_wait': [vp, c_int_p],
        'lbg_pk_prepare_hf': [vp, i, c_double_p, c_double_p],
        'lbg_upload_pk_linear2': [vp, i, c_double_p, c_double_p, c_double_p, c_double_p],
        'lbg_pk_upload_rows': [vp, i, c_double_p, i, c_double_p],
        'lbg_pk_eval': [vp, i, i, c_double_p, i, c_double_p],
        'lbg_eh_transfer': [vp, d, d, d, i, d, d, d, i, c_double_p, c_double_p, c_double_p, c_double_p],
        'lbg_eh_linear_table': [vp, d, d, d, i, d, d, d, d, i, c_double_p, c_double_p, c_double_p],
        'lbg_limber': [vp, i, i, c_double_p, i, c_double_p, c_double_p, c_int_p, i, c_double_p],
        'lbg_forecast_spectra': [vp, i, c_double_p, d, d, d,
                                 i, c_double_p, c_double_p, c_double_p, c_double_p, d, i,
                                 i, c_double_p, c_double_p, c_double_p, c_double_p, d, i,
                                 i, c_double_p, c_double_p, i,
                                 i, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p,
                                 c_double_p, c_double_p, c_double_p],
        'lbg_n0_modi': [vp, i, i, c_double_p, i, c_double_p, c_double_p, i, c_double_p, c_double_p, c_double_p,
                        c_int_p, c_double_p, c_double_p, c_double_p],
        'lbg_eval_spline_ext': [vp, i, d, c_double_p, i, c_double_p, c_double_p],
        'lbg_snr': [vp, i, c_double_p, c_double_p, c_double_p, c_double_p, d, c_double_p],
        'lbg_snr_noalloc': [vp, i, c_double_p, c_double_p, c_double_p, c_double_p, d, c_double_p],
        'lbg_gaussian_cov': [vp, i, c_double_p, i, c_double_p, i, c_int_p, c_int_p, d, c_double_p],
        'lbg_fisher_accum': [vp, i, i, i, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p],
        'lbg_field_fisher': [vp, i, i, i, c_int_p, c_int_p, c_double_p, i, c_double_p, c_double_p, c_double_p, c_double_p],
        'lbg_field_fisher_amp': [vp, i, i, i, c_int_p, c_int_p, c_double_p, i, c_double_p, i, c_double_p, c_double_p, c_double_p, c_double_p,
                                 c_double_p],
        'lbg_limber_keep': [vp, i, i, c_double_p, i, c_double_p, c_double_p, c_int_p, i, i, i, c_int_p],
        'lbg_spectra_set_fill': [vp, i, i, i, c_double_p, i, c_int_p, c_double_p],
        'lbg_spectra_set_get': [vp, i, i, i, i, c_double_p],
        'lbg_field_fisher_sets': [vp, i, i, i, c_int_p, c_int_p, i, i, i, c_int_p, c_int_p, c_double_p, i, c_double_p, c_double_p, c_double_p,
                                  c_double_p],
        'lbg_xchg_create': [vp, i, i, C.c_longlong, C.POINTER(C.c_ubyte)],
        'lbg_xchg_open': [vp, C.POINTER(C.c_ubyte)],
        'lbg_xchg_allgather': [vp, C.c_longlong, c_double_p, c_double_p, c_double_p],
        'lbg_xchg_sums_curves': [vp, i, c_double_p, i, C.c_longlong, c_double_p, c_double_p, c_double_p, c_double_p],
        'lbg_xchg_destroy': [vp],
        'lbg_snr_grid': [vp, i, c_double_p, c_double_p, c_double_p, c_double_p, c_double_p, i, c_double_p, c_double_p, i, c_double_p, i, c_double_p, d,
                         c_double_p],
        'lbg_mqw_fisher': [vp, i, i, c_double_p, i, c_double_p, c_double_p, d, c_double_p, c_double_p, c_double_p],
        'lbg_cadence_fisher': [vp, i, i, c_double_p, c_double_p, c_double_p, d, c_double_p, c_double_p, c_double_p],
        'lbg_forecast_grid': [vp, i, c_double_p, c_double_p, c_double_p, i, c_double_p, c_double_p, c_double_p, c_double_p,
                              c_double_p, i, c_double_p, d, d, c_double_p],
        'lbg_n0_ho02': [vp, i, c_double_p, u, d, d, d, i, i, c_double_p, c_double_p, i, c_double_p, c_double_p,
                        c_double_p, d, d, d, i, i, c_int_p, c_double_p, c_double_p, c_double_p, c_ll_p],
        'lbg_n0_ho02_begin': [vp, i, c_double_p, u, d, d, d, i, i, c_double_p, c_double_p, i, c_double_p, c_double_p,
                              c_double_p, d, d, d, i, i, c_int_p, c_double_p],
        'lbg_n0_ho02_end': [vp, i, c_double_p, c_double_p, c_ll_p, c_ll_p],
        'lbg_n0_ho02_levels': [vp, i, c_double_p, d, i, c_double_p, c_int_p, d, i, c_double_p, c_double_p, i, c_double_p, c_double_p,
                               c_double_p, d, d, d, c_double_p, c_double_p],
        'lbg_n0_ho02_levels_begin': [vp, i, c_double_p, d, i, c_double_p, c_int_p, d, i, c_double_p, c_double_p, i, c_double_p, c_double_p,
                                     c_double_p, d, d, d],
        'lbg_n0_ho02_levels_end': [vp, i, i, c_double_p, c_double_p],
        'lbg_n0_set_exact_below': [vp, d],
        'lbg_n0_last_path': [vp, c_int_p],
        'lbg_scipy_cubic_eval_host': [i, c_double_p, i, c_double_p, c_double_p],
    }
    for name, args in sig.items():
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = C.c_int


def check(rc):
    if rc != 0:
        msg = lib().lbg_last_error().decode('utf-8', 'replace')
        if rc == -1:
            raise ValueError(msg)
        raise LbgError('liblbgcmb error %d: %s' % (rc, msg))


class Context:
    """One per GPU: owns the stream, the uploaded tables and the workspaces (lbg_ctx)."""

    def __init__(self, device=0, lane=0):
        self._h = C.c_void_p()
        self.device = device
        self.lane = lane
        check(lib().lbg_ctx_create_lane(int(device), int(lane), C.byref(self._h)))
        self._cl_key = None

    def close(self):
        if self._h:
            lib().lbg_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    @property
    def handle(self):
        return self._h

    def last_kernel_ms(self):
        ms, n = C.c_double(), C.c_int()
        check(lib().lbg_last_kernel_ms(self._h, C.byref(ms), C.byref(n)))
        return ms.value, n.value

    def last_main_kernel_ms(self):
        v = C.c_double()
        check(lib().lbg_last_main_kernel_ms(self._h, C.byref(v)))
        return v.value

    def kernel_totals(self, reset=False):
        """(device ms, kernel launches) summed over every library call on this context since the last reset."""
        ms, n = C.c_double(), C.c_longlong()
        check(lib().lbg_kernel_totals(self._h, C.byref(ms), C.byref(n), int(bool(reset))))
        return ms.value, n.value

    def fp64_peak_tflops(self):
        v = C.c_double()
        check(lib().lbg_fp64_peak(self._h, C.byref(v)))
        return v.value


def default_device():
    """LOCAL_RANK under torchrun (one process per GPU), else LBGCMB_DEVICE, else 0."""
    return int(os.environ.get('LBGCMB_DEVICE', os.environ.get('LOCAL_RANK', '0')))


def context(device=None, lane=0):
    """Process-wide context cache keyed by (device index, lane).  Lane 0 is the default context; another lane is a second
    context (own stream, own scratch, own mutex) on the same GPU, for calls that should overlap lane-0 work
    (Nkk.n0_terms_async runs on lane 1)."""
    if device is None:
        device = default_device()
    key = device if lane == 0 else (device, lane)
    with _lock:
        ctx = _ctxs.get(key)
    if ctx is None:
        ctx = Context(device, lane)
        with _lock:
            ctx = _ctxs.setdefault(key, ctx)
    return ctx
