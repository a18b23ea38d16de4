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
    """Composite Simpson weights for an odd number of uniform samples (SciPy's simps, cmb/Nkk.py:166-167)."""
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


_DEFAULT_GRIDS = None


def _default_grids():
    """root_grids() and cos / sin of its phi nodes, computed once (the reference rebuilds them per call, cmb/Nkk.py:54-62)."""
    global _DEFAULT_GRIDS
    if _DEFAULT_GRIDS is None:
        ell, w_ell, pphi, w_phi = root_grids()
        _DEFAULT_GRIDS = tuple(_ffi.f64(a) for a in (ell, w_ell, pphi, w_phi, np.cos(pphi), np.sin(pphi)))
        for a in _DEFAULT_GRIDS:
            a.setflags(write=False)
    return _DEFAULT_GRIDS


def _cl_tables(lensCl_interps, nolensCl_interps):
    """(n, lensed [4, n], unlensed [4, n]): the callables sampled on the integer multipoles 0..n-1 the library splines."""
    lens = [table_of(lensCl_interps[x]) for x in _ORDER]
    n = min(t.size for t in lens)
    nolens = []
    for x in _ORDER:
        nolens.append(table_of(nolensCl_interps[x], n_default=n))
    n = min(n, min(t.size for t in nolens))
    return n, _rows_as_matrix(lens, n), _rows_as_matrix(nolens, n)


def _rows_as_matrix(tabs, n):
    """[4, n] float64 C-contiguous matrix of the four tables: the array they are rows of when they already lie that way in memory
    (prep_camb.cl_interps_from_tables builds them from one [4, n] array), a stacked copy otherwise."""
    b = tabs[0].base
    if (isinstance(b, np.ndarray) and b.ndim == 2 and b.shape == (4, n) and b.dtype == np.float64 and b.flags.c_contiguous
            and all(t.base is b and t.shape == (n,) and t.strides == (8,)
                    and t.__array_interface__['data'][0] == b.__array_interface__['data'][0] + i * b.strides[0] for i, t in enumerate(tabs))):
        return b
    return _ffi.f64(np.stack([t[:n] for t in tabs]))


def upload_cls(ctx, lensCl_interps, nolensCl_interps, reuse=False):
    """C_l tables -> the context.  The library compares them, value for value, with the tables it already holds and keeps its
    splines when nothing changed (a sweep over beam / noise calls N0 many times on the same spectra), so every call hands the
    tables over; `reuse` is accepted for compatibility and changes nothing."""
    n, lt, ut = _cl_tables(lensCl_interps, nolensCl_interps)
    _ffi.check(_ffi.lib().lbg_upload_cl_tables(ctx.handle, n, _ffi._dp(lt), _ffi._dp(ut)))
    return n


class N0Call(object):
    """An N0 call in flight (lbg_n0_ho02_begin issued): `result()` waits for it and returns what n0_terms returns."""

    def __init__(self, ctx, n_L, keep):
        self.ctx, self.n_L, self._keep, self._res = ctx, n_L, keep, None

    def result(self):
        if self._res is None:
            import ctypes as C
            out = np.empty(self.n_L)
            out_terms = np.empty((5, self.n_L))
            nodes, unmasked, path = C.c_longlong(0), C.c_longlong(0), C.c_int(0)
            _ffi.check(_ffi.lib().lbg_n0_ho02_end(self.ctx.handle, self.n_L, _ffi._dp(out), _ffi._dp(out_terms), C.byref(nodes),
                                                  C.byref(unmasked)))
            _ffi.check(_ffi.lib().lbg_n0_last_path(self.ctx.handle, C.byref(path)))
            ms, launches = self.ctx.last_kernel_ms()
            last_stats.update(nodes=nodes.value, nodes_unmasked=unmasked.value, kernel_ms=ms, launches=launches, n_L=self.n_L,
                              main_kernel_ms=self.ctx.last_main_kernel_ms(), path=path.value)
            self._keep = None
            self._res = (out, {t: out_terms[i] for i, t in enumerate(('TT', 'TE', 'EE', 'EB', 'BB'))})
        return self._res


def n0_begin(lensCl_interps, nolensCl_interps, Ls, terms=('TT', 'TE', 'EE', 'EB'), thetab=1., DeltaT=1.,
             iterative=False, fold_phi=True, device=None, Tcmb=None, grids=None,
             lmax_T=3000., lmax_P=5000., lmin_v=2., cos_sin=None, exact_u_below=2.0, lane=0, reuse_tables=False,
             exact_rows=None, exact_u=None):
    """Start an N0 evaluation (tables compared / uploaded, inputs staged, kernels enqueued) and return at once: N0Call.

    fold_phi: True = one node per mirror pair (phi, 2 pi - phi), False = every node, 2 = mirror pairs through the generic
    per-node-mask kernel (cross-check).  exact_u_below: l1 rows below it get their u-side spline values from the library's
    SciPy-order host routine (csrc/ctx.cu; DESIGN.md "parity notes"); exact_rows / exact_u let a caller supply its own."""
    for t in terms:
        if t == 'TB':
            raise KeyError('TB')   # the reference indexes nolensCl_interps['TB'], which get_Cls never builds
        if t not in TERM_BITS:
            raise ValueError('unknown estimator %s' % t)
    mask = 0
    for t in terms:
        mask |= TERM_BITS[t]
    ctx = _ffi.context(device, lane)
    upload_cls(ctx, lensCl_interps, nolensCl_interps)
    if grids is None and cos_sin is None:
        ell, w_ell, pphi, w_phi, cphi, sphi = _default_grids()
    else:
        ell, w_ell, pphi, w_phi = grids if grids is not None else root_grids()
        if cos_sin is None:
            cos_sin = (np.cos(pphi), np.sin(pphi))
        cphi, sphi = _ffi.f64(cos_sin[0]), _ffi.f64(cos_sin[1])
        ell, w_ell, w_phi = _ffi.f64(ell), _ffi.f64(w_ell), _ffi.f64(w_phi)
    Ld = _ffi.f64(Ls)
    if Tcmb is None:
        Tcmb = get_params()['Tcmb']
    if exact_rows is None:
        exact_rows, exact_u = np.zeros(0, dtype=np.int32), np.zeros(0)
    exact_rows = np.ascontiguousarray(exact_rows, dtype=np.int32)
    exact_u = _ffi.f64(exact_u)
    fold = 2 if fold_phi == 2 and fold_phi is not True else int(bool(fold_phi))
    _ffi.check(_ffi.lib().lbg_n0_set_exact_below(ctx.handle, float(exact_u_below)))
    _ffi.check(_ffi.lib().lbg_n0_ho02_begin(ctx.handle, Ld.size, _ffi._dp(Ld), mask, float(thetab), float(DeltaT),
                                            float(Tcmb), int(bool(iterative)), ell.size, _ffi._dp(ell), _ffi._dp(w_ell),
                                            pphi.size, _ffi._dp(cphi), _ffi._dp(sphi), _ffi._dp(w_phi),
                                            float(lmax_T), float(lmax_P), float(lmin_v), fold,
                                            exact_rows.size, exact_rows.ctypes.data_as(_ffi.c_int_p), _ffi._dp(exact_u)))
    return N0Call(ctx, Ld.size, (Ld, ell, w_ell, cphi, sphi, w_phi, exact_rows, exact_u))


class N0LevelsCall(object):
    """lbg_n0_ho02_levels_begin issued: `result()` waits and returns (N0[n_lev, n_L], {term: integral[n_lev, n_L]})."""

    def __init__(self, ctx, n_lev, n_L, keep):
        self.ctx, self.n_lev, self.n_L, self._keep, self._res = ctx, n_lev, n_L, keep, None

    def result(self):
        if self._res is None:
            import ctypes as C
            out = np.empty((self.n_lev, self.n_L))
            out_terms = np.empty((self.n_lev, 5, self.n_L))
            _ffi.check(_ffi.lib().lbg_n0_ho02_levels_end(self.ctx.handle, self.n_lev, self.n_L, _ffi._dp(out), _ffi._dp(out_terms)))
            path = C.c_int(0)
            _ffi.check(_ffi.lib().lbg_n0_last_path(self.ctx.handle, C.byref(path)))
            ms, launches = self.ctx.last_kernel_ms()
            last_stats.update(kernel_ms=ms, launches=launches, n_L=self.n_L, n_lev=self.n_lev, main_kernel_ms=self.ctx.last_main_kernel_ms(),
                              path=path.value)
            self._keep = None
            self._res = (out, {t: out_terms[:, i] for i, t in enumerate(('TT', 'TE', 'EE', 'EB', 'BB'))})
        return self._res


def n0_levels_begin(lensCl_interps, nolensCl_interps, Ls, thetab, DeltaTs, iteratives=None, device=None, Tcmb=None, lane=0,
                    lmax_T=3000., lmax_P=5000., lmin_v=2.):
    """Start N0 (TT, TE, EE, EB) for several noise levels DeltaTs[n_lev] of ONE beam `thetab` in one pass of the integrand
    (lbg_n0_ho02_levels: the levels share every table read of a node); `iteratives[n_lev]` as Nkk's `iterative` per level.
    The reference evaluates one experiment at a time (snr.py:108-124 loops cmb/Nkk.py:31 over bolometers)."""
    ctx = _ffi.context(device, lane)
    upload_cls(ctx, lensCl_interps, nolensCl_interps)
    ell, w_ell, pphi, w_phi, cphi, sphi = _default_grids()
    Ld = _ffi.f64(Ls)
    dts = _ffi.f64(np.atleast_1d(DeltaTs))
    its = np.zeros(dts.size, dtype=np.int32) if iteratives is None else np.ascontiguousarray(np.atleast_1d(iteratives), dtype=np.int32)
    if its.size != dts.size:
        raise ValueError('DeltaTs and iteratives differ in length')
    if Tcmb is None:
        Tcmb = get_params()['Tcmb']
    _ffi.check(_ffi.lib().lbg_n0_ho02_levels_begin(ctx.handle, Ld.size, _ffi._dp(Ld), float(thetab), dts.size, _ffi._dp(dts),
                                                   its.ctypes.data_as(_ffi.c_int_p), float(Tcmb), ell.size, _ffi._dp(ell), _ffi._dp(w_ell),
                                                   pphi.size, _ffi._dp(cphi), _ffi._dp(sphi), _ffi._dp(w_phi),
                                                   float(lmax_T), float(lmax_P), float(lmin_v)))
    return N0LevelsCall(ctx, dts.size, Ld.size, (Ld, dts, its))


def n0_levels(*args, **kw):
    """(N0[n_lev, n_L], {term: integral[n_lev, n_L]}) -- arguments of n0_levels_begin."""
    return n0_levels_begin(*args, **kw).result()


def n0_terms(*args, **kw):
    """(N0[len(Ls)], {term: integral[len(Ls)]}) -- the combined noise plus the per-estimator integrals (arguments of n0_begin)."""
    return n0_begin(*args, **kw).result()


def n0_terms_async(*args, **kw):
    """`n0_begin` on a second context of the same GPU (lane 1: own stream, scratch and phi-table block, lowest stream priority);
    returns the N0Call, whose result() is what n0_terms returns.

    The N0 curve is the long pole of a forecast (25 ms for 4951 L) and nothing in the Limber stage depends on it until the
    variances are formed (Ckg.py:83-87), so a driver starts it first and builds its spectra meanwhile: one host thread, the two
    contexts' kernels share the GPU (the short lane-0 kernels take SM slots as N0 CTAs retire)."""
    kw.setdefault('lane', 1)
    return n0_begin(*args, **kw)


class _Mapped(object):
    def __init__(self, call, fn):
        self._call, self._fn = call, fn

    def result(self):
        return self._fn(self._call.result())


def Nkk_async(lensCl_interps, nolensCl_interps, Ls, terms=['TT', 'TE', 'EE', 'EB'], thetab=1., DeltaT=1., iterative=False):
    """Nkk(...) in flight on lane 1 (no pickle cache): result() returns the curve.  See n0_terms_async."""
    return _Mapped(n0_begin(lensCl_interps, nolensCl_interps, Ls, terms, thetab, DeltaT, iterative, lane=1), lambda r: r[0])


def _cache_path(Ls, terms, thetab, DeltaT, iterative, lensCl_interps=None, nolensCl_interps=None):
    """The reference's file name (cmb/Nkk.py:36-41) + a digest of what that name leaves out: the C_l tables (hence the cosmology) and
    the exact L grid.  Under the reference's key a file written for other spectra is silently served; here it is simply another file."""
    root = os.environ['LBGCMB']
    fpath = root + '/pickle/Nkk_thetab_%.2lf_DeltaT_%.2lf_' % (thetab, DeltaT) + 'Nlls_%d_Lmin_%d_Lmax_%d_' % (len(Ls), Ls[0], Ls[-1])
    fpath += '_'.join(terms) + '_iterative_%d' % int(iterative == True)  # noqa: E712 (reference spelling)
    if lensCl_interps is not None:
        import hashlib
        h = hashlib.sha256()
        _, lt, ut = _cl_tables(lensCl_interps, nolensCl_interps)
        for a in (lt, ut, _ffi.f64(np.asarray(Ls, dtype=np.float64)), np.array([thetab, DeltaT], dtype=np.float64)):
            h.update(a.tobytes())
        fpath += '_cl_%s' % h.hexdigest()[:16]
    return fpath + '.p'


def Nkk(lensCl_interps, nolensCl_interps, Ls, terms=['TT', 'TE', 'EE', 'EB'], thetab=1., DeltaT=1., iterative=False,
        pickleit=True):
    """Ckk noise curve for the flat-sky limit (Hu & Okamoto 2002); drop-in for cmb/Nkk.py:31.

    The pickle cache of cmb/Nkk.py:36-46,174-182 is honoured only when $LBGCMB is set; its file name carries a
    digest of the C_l tables and the L grid, which the reference's key ignores."""
    have_root = 'LBGCMB' in os.environ
    if have_root:
        try:
            return load(open(_cache_path(Ls, terms, thetab, DeltaT, iterative, lensCl_interps, nolensCl_interps), 'rb'))
        except Exception:
            pass
    result, _ = n0_terms(lensCl_interps, nolensCl_interps, Ls, terms, thetab, DeltaT, iterative)
    if pickleit and have_root:
        fpath = _cache_path(Ls, terms, thetab, DeltaT, iterative, lensCl_interps, nolensCl_interps)
        os.makedirs(os.path.dirname(fpath), exist_ok=True)
        dump(result, open(fpath, 'wb'))
    return result
