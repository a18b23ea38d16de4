"""Hu-Okamoto N0 lensing-reconstruction noise; mirrors reference cmb/Nkk.py:31-184 (`Nkk`).

Same signature, argument meaning, defaults and return type (float64[len(Ls)]).  The (phi, l1, L) cube of the
reference (cmb/Nkk.py:54-167) is never materialised: the quadrature nodes/weights are built here exactly as
the reference builds them and the integrand is evaluated by the sm_100a kernel in csrc/n0.cu through
lbg_n0_ho02.  There is no CPU path.
"""
import os
from pickle import load, dump

import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200.params import get_params
from lbgcmb_b200.prep_camb import table_of

TERM_BITS = {'TT': 1, 'TE': 2, 'EE': 4, 'EB': 8, 'BB': 16}
_ORDER = ('TT', 'EE', 'BB', 'TE')

# set by every Nkk()/n0_terms() call: integrand nodes executed, kernel ms, launches
last_stats = {}


def prep_lmaxmask(nu, nv):
    """The multipole cuts of cmb/Nkk.py:13-26 as 0/1 arrays per estimator: |u| and |v| at most 3000 for the temperature
    estimators and 5000 for the polarisation ones, |v| at least 2.  Host helper with the reference's return type; the kernel
    applies the same cuts as phi-index ranges (csrc/n0.cu) and never forms these arrays."""
    nu, nv = np.asarray(nu), np.asarray(nv)
    cut = {}
    for lmax in (3000, 5000):
        keep = np.ones_like(nu) if lmax == 3000 else np.ones_like(nv)
        keep[(nu > lmax) | (nv > lmax) | (nv < 2)] = 0.0
        cut[lmax] = keep
    return {'TT': cut[3000], 'TE': cut[3000], 'EE': cut[5000], 'BB': cut[5000], 'TB': cut[5000], 'EB': cut[5000]}


def _simpson_weights(n, dx):
    """Composite Simpson weights for an odd number of uniform samples (scipy simps, cmb/Nkk.py:166-167)."""
    if n % 2 == 0:
        raise ValueError('the reference N0 grids are odd-length (SURVEY.md N5)')
    w = np.zeros(n)
    w[0:n - 2:2] += dx / 3.0
    w[1:n - 1:2] += 4.0 * dx / 3.0
    w[2:n:2] += dx / 3.0
    return w


def root_grids(lmax=5000., Nphi=200, Nell=10000):
    """Nodes and weights of cmb/Nkk.py:54-62,166-167: ell (9999), phi (201); note the swapped dx labels (N5)."""
    dphi = 2. * np.pi / Nphi
    dell = lmax / Nell
    ell = np.arange(1.0, lmax + dell, dell)
    pphi = np.arange(0.0, 2. * np.pi + dphi, dphi)
    w_phi = _simpson_weights(pphi.size, dell)     # first simps pass runs over phi with dx=dell
    w_ell = _simpson_weights(ell.size, dphi)      # second over l1 with dx=dphi
    return ell, w_ell, pphi, w_phi


def upload_cls(ctx, lensCl_interps, nolensCl_interps, reuse=False):
    """C_l tables -> the context (spline solve + H2D).  `reuse`: skip the upload when the tables equal, value for value, the ones
    this context received last (a sweep over beam / noise calls N0 many times on the same spectra)."""
    lens = [table_of(lensCl_interps[x]) for x in _ORDER]
    n = min(t.size for t in lens)
    nolens = []
    for x in _ORDER:
        nolens.append(table_of(nolensCl_interps[x], n_default=n))
    n = min(n, min(t.size for t in nolens))
    lt = _ffi.f64(np.stack([t[:n] for t in lens]))
    ut = _ffi.f64(np.stack([t[:n] for t in nolens]))
    last = getattr(ctx, '_cl_last', None)
    same = last is not None and last[0].shape == lt.shape and np.array_equal(last[0], lt) and np.array_equal(last[1], ut)
    if not same:
        ctx._exact_cache = None              # host-side values derived from the tables (_exact_u_rows)
    if same and reuse:
        return n
    _ffi.check(_ffi.lib().lbg_upload_cl_tables(ctx.handle, n, _ffi._dp(lt), _ffi._dp(ut)))
    ctx._cl_last = (lt, ut)
    return n


def _exact_u_rows(lensCl_interps, nolensCl_interps, ell, cphi, sphi, below):
    """u-side spline values at |u|_j for the l1 rows below the first non-zero C_l (l1 < 2: CAMB tables have C_0 = C_1 = 0).

    There C^tot(|u|) is pure detector noise (~1e-20) while the spline crosses zero between knots of size 1e-10, so the
    reference's 1/C^tot carries the last-bit rounding of scipy's `interp1d` evaluation at relative 1e-6 -- and these rows
    carry up to 20 % of the TT integral at low L.  For bit-faithful parity the few affected values (2 rows x 201 nodes x 8
    spectra) are evaluated with the very routine the reference uses (prep_camb.py:88-89) and handed to the kernel
    (lbg_n0_ho02 `exact_u`); every other node is evaluated on the device."""
    from scipy.interpolate import interp1d
    rows = np.nonzero(ell < below)[0].astype(np.int32)
    if rows.size == 0:
        return rows, np.zeros(0)
    ux = ell[rows][:, None] * cphi[None, :]
    uy = ell[rows][:, None] * sphi[None, :]
    nu = (ux ** 2. + uy ** 2.) ** 0.5                          # cmb/Nkk.py:79
    vals = np.empty((rows.size, cphi.size, 8))
    k = 0
    for interps in (nolensCl_interps, lensCl_interps):
        for name in ('TT', 'TE', 'EE', 'BB'):
            f = interps[name]
            if not isinstance(f, interp1d):
                g = getattr(f, '_scipy_interp1d', None)                # built once per table-backed callable
                if g is None:
                    g = interp1d(np.arange(table_of(f).size, dtype=np.float64), table_of(f), kind='cubic',
                                 bounds_error=False, fill_value=0.0)
                    try:
                        f._scipy_interp1d = g
                    except AttributeError:
                        pass
                f = g
            vals[:, :, k] = f(nu)
            k += 1
    return np.ascontiguousarray(rows), np.ascontiguousarray(vals)


def n0_terms(lensCl_interps, nolensCl_interps, Ls, terms=('TT', 'TE', 'EE', 'EB'), thetab=1., DeltaT=1.,
             iterative=False, fold_phi=True, device=None, Tcmb=None, grids=None,
             lmax_T=3000., lmax_P=5000., lmin_v=2., cos_sin=None, exact_u_below=2.0, lane=0, reuse_tables=False):
    """(N0[len(Ls)], {term: integral[len(Ls)]}) -- the combined noise plus the per-estimator integrals."""
    for t in terms:
        if t == 'TB':
            raise KeyError('TB')   # the reference indexes nolensCl_interps['TB'], which get_Cls never builds
        if t not in TERM_BITS:
            raise ValueError('unknown estimator %s' % t)
    mask = 0
    for t in terms:
        mask |= TERM_BITS[t]
    ctx = _ffi.context(device, lane)
    upload_cls(ctx, lensCl_interps, nolensCl_interps, reuse=reuse_tables)
    ell, w_ell, pphi, w_phi = grids if grids is not None else root_grids()
    Ld = _ffi.f64(Ls)
    if cos_sin is None:
        cos_sin = (np.cos(pphi), np.sin(pphi))
    cphi, sphi = _ffi.f64(cos_sin[0]), _ffi.f64(cos_sin[1])
    ell, w_ell, w_phi = _ffi.f64(ell), _ffi.f64(w_ell), _ffi.f64(w_phi)
    out = np.empty(Ld.size)
    out_terms = np.empty((5, Ld.size))
    import ctypes as C
    # the scipy-evaluated u-side rows depend on the tables and the grids only: kept while the context's tables are value-equal
    ekey = (ell.size, float(exact_u_below), cphi.tobytes(), sphi.tobytes(), ell.tobytes())
    cached = getattr(ctx, '_exact_cache', None)
    if cached is not None and cached[0] == ekey:
        exact_rows, exact_u = cached[1], cached[2]
    else:
        exact_rows, exact_u = _exact_u_rows(lensCl_interps, nolensCl_interps, ell, cphi, sphi, exact_u_below)
        ctx._exact_cache = (ekey, exact_rows, exact_u)
    nodes = C.c_longlong(0)
    if Tcmb is None:
        Tcmb = get_params()['Tcmb']
    _ffi.check(_ffi.lib().lbg_n0_ho02(ctx.handle, Ld.size, _ffi._dp(Ld), mask, float(thetab), float(DeltaT),
                                      float(Tcmb), int(bool(iterative)), ell.size, _ffi._dp(ell), _ffi._dp(w_ell),
                                      pphi.size, _ffi._dp(cphi), _ffi._dp(sphi), _ffi._dp(w_phi),
                                      float(lmax_T), float(lmax_P), float(lmin_v), int(bool(fold_phi)),
                                      exact_rows.size, exact_rows.ctypes.data_as(_ffi.c_int_p), _ffi._dp(exact_u),
                                      _ffi._dp(out), _ffi._dp(out_terms), C.byref(nodes)))
    ms, launches = ctx.last_kernel_ms()
    last_stats.update(nodes=nodes.value, kernel_ms=ms, launches=launches, n_L=Ld.size, main_kernel_ms=ctx.last_main_kernel_ms())
    return out, {t: out_terms[i] for i, t in enumerate(('TT', 'TE', 'EE', 'EB', 'BB'))}


_pool = None


def _submit(fn):
    global _pool
    if _pool is None:
        from concurrent.futures import ThreadPoolExecutor
        _pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix='lbg-n0')
    return _pool.submit(fn)


def n0_terms_async(*args, **kw):
    """`n0_terms` submitted to a worker thread on a second context of the same GPU (lane 1: own stream and scratch);
    returns a concurrent.futures.Future whose result() is what n0_terms returns.

    The N0 curve is the long pole of a forecast (25 ms for 4951 L) and nothing in the Limber stage depends on it until
    the variances are formed (Ckg.py:83-87), so a driver can start it first and build its spectra meanwhile: the library
    call releases the GIL and the two contexts' kernels share the GPU."""
    kw.setdefault('lane', 1)
    return _submit(lambda: n0_terms(*args, **kw))


def Nkk_async(lensCl_interps, nolensCl_interps, Ls, terms=['TT', 'TE', 'EE', 'EB'], thetab=1., DeltaT=1., iterative=False):
    """Future of Nkk(...) (no pickle cache): see n0_terms_async."""
    return _submit(lambda: n0_terms(lensCl_interps, nolensCl_interps, Ls, terms, thetab, DeltaT, iterative, lane=1)[0])


def _cache_path(Ls, terms, thetab, DeltaT, iterative):
    root = os.environ['LBGCMB']
    fpath = root + '/pickle/Nkk_thetab_%.2lf_DeltaT_%.2lf_' % (thetab, DeltaT) + 'Nlls_%d_Lmin_%d_Lmax_%d_' % (len(Ls), Ls[0], Ls[-1])
    fpath += '_'.join(terms) + '_iterative_%d.p' % int(iterative == True)  # noqa: E712 (reference spelling)
    return fpath


def Nkk(lensCl_interps, nolensCl_interps, Ls, terms=['TT', 'TE', 'EE', 'EB'], thetab=1., DeltaT=1., iterative=False,
        pickleit=True):
    """Ckk noise curve for the flat-sky limit (Hu & Okamoto 2002); drop-in for cmb/Nkk.py:31.

    The pickle cache of cmb/Nkk.py:36-46,174-182 is honoured only when $LBGCMB is set."""
    have_root = 'LBGCMB' in os.environ
    if have_root:
        try:
            return load(open(_cache_path(Ls, terms, thetab, DeltaT, iterative), 'rb'))
        except Exception:
            pass
    result, _ = n0_terms(lensCl_interps, nolensCl_interps, Ls, terms, thetab, DeltaT, iterative)
    if pickleit and have_root:
        fpath = _cache_path(Ls, terms, thetab, DeltaT, iterative)
        os.makedirs(os.path.dirname(fpath), exist_ok=True)
        dump(result, open(fpath, 'wb'))
    return result
