"""Background + P(k,z); mirrors reference modi/cosmology.py:17-194 (`Cosmology`).

Host side (as in the reference): E(a), chi(a) by `quad` (:134-150), un-normalised D+ on 1000 ln a nodes + cubic spline
(:92-117; the reference integrates with scipy-1.2.1 `romberg`, since removed from scipy and restated in
lbgcmb_b200/quadrature.py, so D+ matches the reference's own value rather than the exact integral).
Device side (csrc/limber.cu): pkalin / pkanlin rows for any number of redshifts at once (`pk_rows`), HALOFIT included.
"""
import numpy as np
from scipy.integrate import quad
from scipy.interpolate import InterpolatedUnivariateSpline as interpolate

from lbgcmb_b200 import _ffi
from lbgcmb_b200.cosmo import FlatLambdaCDM
from lbgcmb_b200.modi import halofit as _hf
from lbgcmb_b200.quadrature import simps_weights_x, romberg

_owner = _ffi.table_owner
_PK_SLOTS = 4                            # lbg_ctx::PK_SLOTS
# the grid modi/halofit.py:36-37 fits ln sigma^2 on, and its exp, as numpy computes them
_HF_LNR = _ffi.f64(np.linspace(np.log(0.1), np.log(10.0), 500))
_HF_R = _ffi.f64(np.exp(_HF_LNR))


LOGA_NODES = np.linspace(-20., 0, 1000)            # modi/cosmology.py:98-101


def growth_tables(Ms, Ls=None, device=None, want_levels=False):
    """Un-normalised D+ on the reference's 1000 ln a nodes for many cosmologies in one launch (csrc/growth.cu): row c is the `y`
    of modi/cosmology.py:107 for (M, L) = (Ms[c], Ls[c]) (L defaults to 1 - M).  Pass a row to Cosmology(Dplus_table=...)."""
    Ms = _ffi.f64(np.atleast_1d(Ms))
    Ls = _ffi.f64(1. - Ms if Ls is None else np.atleast_1d(Ls))
    if Ms.shape != Ls.shape:
        raise ValueError('Ms and Ls differ in shape')
    import ctypes as C
    out = np.empty((Ms.size, LOGA_NODES.size))
    lev = np.empty((Ms.size, LOGA_NODES.size), dtype=np.int32) if want_levels else None
    ctx = _ffi.context(device)
    _ffi.check(_ffi.lib().lbg_growth_table(ctx.handle, Ms.size, _ffi._dp(Ms), _ffi._dp(Ls), LOGA_NODES.size, _ffi._dp(LOGA_NODES),
                                           _ffi._dp(out), lev.ctypes.data_as(C.POINTER(C.c_int)) if want_levels else None))
    return (out, lev) if want_levels else out


class Cosmology(object):
    def __init__(self, M=0.292, L=None, H0=100., B=None, sig8=0.819, klin=None, plin=None, pfile=None, device=None,
                 Dplus_table=None):
        self.M = M
        if L is None:
            L = 1 - M
        self.L = L
        self.B = B
        self.H0 = H0
        self.sig8 = sig8
        self.cin = 1000. / (299792458.)
        self.cosmo = FlatLambdaCDM(H0=H0, Om0=M, Ob0=B, name="cosmo")
        self.pfile = pfile
        if pfile is not None:
            self.klin, self.plin = np.loadtxt(pfile, unpack=True)
        else:
            self.klin, self.plin = klin, plin
        # Dplus_table: the 1000 un-normalised D+ values from growth_tables() (one launch for a whole cosmology sweep) instead of
        # 1000 host Romberg integrals (0.11 s per cosmology); 'device' builds this one cosmology's table on the GPU
        if Dplus_table is None:
            self.Dplus = self._eval_Dplus()
        else:
            y = growth_tables([M], [L], device=device)[0] if isinstance(Dplus_table, str) else np.asarray(Dplus_table, dtype=np.float64)
            self.Dplus = interpolate(np.exp(LOGA_NODES), y)
        self.device = device
        self._token = _ffi.new_owner_token()      # identifies this object's table on a context (id() is reused by CPython)
        self._wlnk = None
        self._sig0 = None
        self._spec_cache = {}
        self._async_keys = None
        self._bg_cache = {}              # per z grid: D(z)^2, Omega_m(z), Omega_L(z) -- the background, independent of the P(k) table
        self._rows_key = None
        self._slots = {}                 # z-grid key -> device table slot (lbg_pk_select_slot); insertion order = LRU order
        self._cur_slot = 0

    # ---- background (host) ------------------------------------------------------------------------------------
    def _za(self, z, a):
        if a is None:
            if z is None:
                print("At least one of scale factor or redshift is needed.")
                return None
            return z, self.ztoa(z)
        return self.atoz(a), a

    def atoz(self, a):
        return 1. / a - 1.

    def ztoa(self, z):
        return 1. / (1. + z)

    def OmegaMa(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.M / a ** 3. / (self.M / a ** 3. + self.L)

    def Ea(self, z=None, a=None):
        z, a = self._za(z, a)
        M, L = self.M, self.L
        return a ** -1.5 * (M + (1. - M - L) * a + L * a ** 3.) ** 0.5

    def Ha(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.H0 * self.Ea(a=a)

    def _eval_Dplus(self):
        """Un-normalised D(a) on 1000 ln a nodes: Romberg from ln a = -20 to each node (modi/cosmology.py:92-111)."""
        logx = np.linspace(-20., 0, 1000)
        x = np.exp(logx)

        def kernel(loga):
            a = np.exp(loga)
            return (a * self.Ea(a=a)) ** -3. * a
        y = self.Ea(a=x) * np.array([romberg(kernel, logx.min(), loga) for loga in logx])
        return interpolate(x, y)

    def Dgrow(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.Dplus(a) / self.Dplus(1.)

    def chia(self, z=None, a=None):
        """comoving distance to scale factor a in Mpc/h (modi/cosmology.py:134-150)"""
        fac = 100. * self.cin
        f = lambda x: (self.Ea(a=x) * x ** 2) ** -1.
        z, a = self._za(z, a)
        if type(a) == np.ndarray:
            y = np.zeros_like(a)
            for foo in range(len(a)):
                y[foo] = quad(f, a[foo], 1)[0]
            return y / fac
        return quad(f, a, 1)[0] / fac

    def k_chil(self, l, z=None, a=None):
        z, a = self._za(z, a)
        return (l + 0.5) / self.chia(a=a)

    # ---- P(k, z) (device) ---------------------------------------------------------------------------------------
    def _ctx(self):
        ctx = _ffi.context(self.device)
        k, p = _ffi.f64(self.klin), _ffi.f64(self.plin)
        owner = (self._token, _ffi.table_digest(k, p))           # the digest catches tables edited in place
        if _owner.get(ctx.device) != owner:
            if self._wlnk is None or self._wlnk[0].size != k.size or not np.array_equal(self._wlnk[2], k):
                lnk = _ffi.f64(np.log(k))
                self._wlnk = (_ffi.f64(simps_weights_x(lnk)), lnk, k.copy())
            _ffi.check(_ffi.lib().lbg_upload_pk_linear2(ctx.handle, k.size, _ffi._dp(k), _ffi._dp(p), _ffi._dp(self._wlnk[0]),
                                                        _ffi._dp(self._wlnk[1])))
            s8 = self.sig8
            if s8 < 1.0 and s8 > 0.6:        # the HALOFIT sigma^2 sums start now, unwaited: build_tables picks them up
                _ffi.check(_ffi.lib().lbg_pk_prepare_hf(ctx.handle, _HF_LNR.size, _ffi._dp(_HF_LNR), _ffi._dp(_HF_R)))
            _owner[ctx.device] = owner
            self._rows_key = None
            self._slots = {}             # the upload dropped every P(k,z) table of the context
            self._sig0 = None
            self._spec_cache = {}
        return ctx

    def invalidate_device(self):
        """Forget what this object has on the GPU: the next call uploads the linear table and rebuilds the P(k,z) rows."""
        self.finish_tables()
        dev = _ffi.context(self.device).device
        if (_owner.get(dev) or (None,))[0] == self._token:
            _owner[dev] = None
        self._rows_key = None
        self._slots = {}

    def _sigma2_z0(self, R):
        ctx = self._ctx()
        R = _ffi.f64(np.atleast_1d(R))
        out = np.empty(R.size)
        _ffi.check(_ffi.lib().lbg_halofit_sigma2(ctx.handle, R.size, _ffi._dp(R), _ffi._dp(out)))
        return out

    def halofit_spec(self, z, sig8=None):
        """(k_nl, n_eff, curvature) at redshift z: modi/halofit.py:7-100 applied to Delta^2_lin(k) D(z)^2 (its own pair of quintic
        fits on the host, exactly as the reference does for every call; FITPACK order, csrc/fitpack.cuh)."""
        sig8 = self.sig8 if sig8 is None else sig8
        self._ctx()
        key = (float(z), float(sig8))
        if key not in self._spec_cache:
            D2 = float(self.Dgrow(z=z)) ** 2.
            if sig8 < 1.0 and sig8 > 0.6:
                if self._sig0 is None:
                    self._sig0 = self._sigma2_z0(_HF_R)
                spec = _hf.spec_from_lnsig(_HF_LNR, np.log(D2 * self._sig0), lambda R: D2 * self._sigma2_z0(R))
            else:
                spec = _hf._get_spec_from(lambda R: D2 * self._sigma2_z0(R), sig8)
            self._spec_cache[key] = spec
        return self._spec_cache[key]

    def finish_tables(self):
        """Wait for a build started with build_tables(..., wait=False).  True: the tables are good; False: some redshift needs the
        reference's exception path -- the tables were dropped, call build_tables(grids) (blocking) and redo what used them."""
        import ctypes as C
        if not self._async_keys:
            return True
        keys, self._async_keys = self._async_keys, None
        n = C.c_int(0)
        _ffi.check(_ffi.lib().lbg_pk_build_wait(_ffi.context(self.device).handle, C.byref(n)))
        if n.value:
            for key in keys:
                self._slots.pop(key, None)
            self._rows_key = None
            return False
        return True

    def _device_tables(self, ctx, zins, slots, build, want_spec=False, out=None):
        """lbg_pk_build_tables_hf on one or more z grids: zins = [rows[n_i, 4] = (D^2, Omega_m(z), Omega_L(z), use_halofit)], slots =
        their table slots (-1: the current one).  Returns (spec[sum n_i, 4] or None, n_flagged)."""
        import ctypes as C
        n_each = np.ascontiguousarray([z.shape[0] for z in zins], dtype=np.int32)
        sl = np.ascontiguousarray(slots, dtype=np.int32)
        zin = _ffi.f64(np.concatenate(zins)) if len(zins) > 1 else _ffi.f64(zins[0])
        spec = np.empty((zin.shape[0], 4)) if want_spec else None
        nflag = C.c_int(0)
        fnu = self.cosmo.Onu0 / self.cosmo.Om0
        _ffi.check(_ffi.lib().lbg_pk_build_tables_hf(ctx.handle, len(zins), n_each.ctypes.data_as(_ffi.c_int_p), sl.ctypes.data_as(_ffi.c_int_p),
                                                     _ffi._dp(zin), _HF_LNR.size, _ffi._dp(_HF_LNR), _ffi._dp(_HF_R), float(fnu),
                                                     float(self.cosmo.Om0), -1.0, 1, int(build), _ffi._dp(spec) if want_spec else None,
                                                     C.byref(nflag), _ffi._dp(out) if out is not None else None))
        return spec, nflag.value

    def halofit_specs(self, zs, D2, sig8=None):
        """halofit_spec for many redshifts at once -> [n, 3] = (k_nl, n_eff, curvature), on the device (csrc/halofit.cu): R_nl(z) from
        one inverse fit at z = 0 evaluated at -ln D^2 (ln sigma^2(R, z) = ln D^2 + ln sigma_0^2(R); 1e-14 against per-z inverse fits),
        n_eff and the curvature from a per-redshift quintic fit in FITPACK order (one device thread per redshift).  Redshifts whose
        R_nl leaves the tabulated decade, and unusual sigma_8, take the per-z host route, which reproduces the reference's other
        branches."""
        sig8 = self.sig8 if sig8 is None else sig8
        zs, D2 = np.asarray(zs, dtype=np.float64), np.asarray(D2, dtype=np.float64)
        out = np.empty((zs.size, 3))
        if not (sig8 < 1.0 and sig8 > 0.6):
            for i, z in enumerate(zs):
                out[i] = self.halofit_spec(z, sig8)
            return out
        self.finish_tables()
        ctx = self._ctx()
        zin = np.zeros((zs.size, 4))
        zin[:, 0] = D2
        zin[:, 3] = 1.0
        spec, _ = self._device_tables(ctx, [_ffi.f64(zin)], [-1], build=False, want_spec=True)
        out[:, 0] = spec[:, 0]
        out[:, 1:3] = spec[:, 1:3]
        for i in np.nonzero(spec[:, 3] != 0.0)[0]:
            out[i] = self.halofit_spec(zs[i], sig8)
        return out

    def build_rows(self, zs, nonlinear=True, sig8=None, zmax_halofit=4.0, want_host=False):
        """Device-resident table with row r = P(k, zs[r]): pkanlin (HALOFIT for z <= 4, else linear) or pkalin."""
        zs = np.atleast_1d(np.asarray(zs, dtype=np.float64))
        key = (zs.tobytes(), bool(nonlinear), sig8, zmax_halofit)
        self.finish_tables()
        ctx = self._ctx()
        if self._rows_key == key and not want_host:
            return None
        slot = self._slots.pop(key, None)
        if slot is not None and not want_host:                  # built earlier for this z grid and still parked on the device
            self._slots[key] = slot
            self._select_slot(ctx, slot)
            self._rows_key = key
            return None
        if slot is None:
            # single-row tables (pk_eval, pkalin / pkanlin) share the last slot so that they do not evict the z-grid tables
            small = zs.size == 1
            pool = (_PK_SLOTS - 1,) if small else range(_PK_SLOTS - 1)
            used = set(self._slots.values())
            free = [i for i in pool if i not in used]
            if free:
                slot = free[0]
            else:
                old = next(k for k, v in self._slots.items() if v in pool)   # least recently used
                slot = self._slots.pop(old)
        self._select_slot(ctx, slot)
        bg = self._background(key, zs, nonlinear, zmax_halofit)
        import ctypes as C
        nk = C.c_int(0)
        _ffi.check(_ffi.lib().lbg_pk_table_nk(ctx.handle, C.byref(nk)))
        if nk.value != len(self.klin):
            raise RuntimeError('the context holds a %d-node k grid, this object has %d' % (nk.value, len(self.klin)))
        out = np.empty((zs.size, len(self.klin))) if want_host else None
        fnu = self.cosmo.Onu0 / self.cosmo.Om0
        std_sig8 = (self.sig8 if sig8 is None else sig8)
        std_sig8 = std_sig8 < 1.0 and std_sig8 > 0.6
        done = False
        if std_sig8 or not np.any(bg[:, 3]):
            # spectral parameters and rows on the device in one call (no host round trip); flagged rows -> explicit route below
            _, nflag = self._device_tables(ctx, [bg], [-1], build=True, out=out)
            done = nflag == 0
        if not done:
            rows = np.zeros((zs.size, 8))
            rows[:, 0] = bg[:, 0]
            rows[:, 1] = 1.0
            hf = bg[:, 3] != 0.0
            if np.any(hf):
                rows[hf, 1:4] = self.halofit_specs(zs[hf], rows[hf, 0], sig8)
                rows[hf, 4] = bg[hf, 1]
                rows[hf, 5] = bg[hf, 2]
                rows[hf, 6] = 1.0
            _ffi.check(_ffi.lib().lbg_pk_build_table(ctx.handle, zs.size, _ffi._dp(_ffi.f64(rows)), float(fnu), float(self.cosmo.Om0),
                                                     -1.0, 1, _ffi._dp(out) if want_host else None))
        self._rows_key = key
        self._slots[key] = slot
        return out

    def build_tables(self, grids, nonlinear=True, sig8=None, zmax_halofit=4.0, wait=True):
        """Table slots for several z grids at once, every missing table built by ONE library call (lbg_pk_build_tables_hf: one
        sigma^2 pass, one spectral-parameter launch for all rows, one row launch per table).  Falls back to table_slot per grid for
        unusual sigma_8 or when a redshift needs the reference's exception path.  wait=False: the build is enqueued and not waited
        for (the caller evaluates its p(z), b(z) callables meanwhile; any library call made afterwards is ordered behind it) and
        finish_tables() must be called before results that used the tables are trusted."""
        grids = [np.atleast_1d(np.asarray(z, dtype=np.float64)) for z in grids]
        keys = [(z.tobytes(), bool(nonlinear), sig8, zmax_halofit) for z in grids]
        s8 = self.sig8 if sig8 is None else sig8
        if len(grids) > _PK_SLOTS - 1 or not (s8 < 1.0 and s8 > 0.6):
            return [self.table_slot(z, nonlinear=nonlinear, sig8=sig8) for z in grids]
        ctx = self._ctx()
        for key in keys:                                   # refresh the recency of the tables that are already parked
            if key in self._slots:
                self._slots[key] = self._slots.pop(key)
        missing = [i for i, key in enumerate(keys) if key not in self._slots]
        if missing:
            pool = range(_PK_SLOTS - 1)
            zins, slots = [], []
            for i in missing:
                used = set(self._slots.values())
                free = [q for q in pool if q not in used]
                if free:
                    slot = free[0]
                else:
                    old = next(k for k, v in self._slots.items() if v in pool and k not in keys)   # least recently used, not of this batch
                    slot = self._slots.pop(old)
                self._slots[keys[i]] = slot
                zins.append(self._background(keys[i], grids[i], nonlinear, zmax_halofit))
                slots.append(slot)
            self.finish_tables()
            _, nflag = self._device_tables(ctx, zins, slots, build=(1 if wait else 2))
            self._rows_key = None                          # whichever slot is current may just have been rewritten: select before use
            if not wait:
                self._async_keys = [keys[i] for i in missing]
                nflag = 0
            if nflag:
                for i in missing:
                    del self._slots[keys[i]]
                return [self.table_slot(z, nonlinear=nonlinear, sig8=sig8) for z in grids]
        return [self._slots[key] for key in keys]

    def _background(self, key, zs, nonlinear, zmax_halofit):
        """Per z grid: D^2, Omega_m(z), Omega_L(z), use_halofit -- the background, kept across table rebuilds."""
        bg = self._bg_cache.get(key)
        if bg is None:
            bg = np.zeros((zs.size, 4))
            aa = self.ztoa(zs)
            bg[:, 0] = (self.Dplus(aa) / self.Dplus(1.)) ** 2.                   # Dgrow(z)^2, modi/cosmology.py:113-117
            if nonlinear:
                hf = ~(zs > zmax_halofit)                                         # modi/cosmology.py:171
                zz = self.atoz(aa[hf])
                bg[hf, 1] = self.cosmo.Om(zz)
                bg[hf, 2] = self.cosmo.Ode(zz)
                bg[hf, 3] = 1.0
            bg = _ffi.f64(bg)
            if len(self._bg_cache) >= 16:
                self._bg_cache.clear()
            self._bg_cache[key] = bg
        return bg

    def table_slot(self, zs, nonlinear=True, sig8=None):
        """Device table slot (lbg_pk_select_slot index) holding row r = P(k, zs[r]), built if it is not parked already: what
        lbg_forecast_spectra takes per z grid."""
        self.build_rows(zs, nonlinear=nonlinear, sig8=sig8)
        return self._cur_slot

    def _select_slot(self, ctx, slot):
        import ctypes as C
        n = C.c_int(0)
        _ffi.check(_ffi.lib().lbg_pk_select_slot(ctx.handle, int(slot), C.byref(n)))
        self._cur_slot = slot

    def pkalin(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.klin, self.build_rows([z], nonlinear=False, want_host=True)[0]

    def pkanlin(self, z=None, a=None, sig8=None):
        z, a = self._za(z, a)
        return self.klin, self.build_rows([z], nonlinear=True, sig8=sig8, want_host=True)[0]

    def limber(self, ells, zs, chis, weights, nonlinear=True, oob=0, sig8=None, keep=None):
        """C_l[b] = sum_z weights[b, z] P(k=(l+1/2)/chi_z, z) on the device (csrc/limber.cu limber_kernel).
        keep = (set, set_rows, dest_rows): the spectra stay on the device, batch row b as row dest_rows[b] of spectra set `set`
        (lbg_limber_keep, for fisher.field_fisher_sets); returns None."""
        zs = np.atleast_1d(np.asarray(zs, dtype=np.float64))
        self.build_rows(zs, nonlinear=nonlinear, sig8=sig8)
        ctx = self._ctx()
        ells, chis = _ffi.f64(np.atleast_1d(ells)), _ffi.f64(np.atleast_1d(chis))
        w = _ffi.f64(np.atleast_2d(weights))
        prow = np.arange(zs.size, dtype=np.int32)
        if chis.size != zs.size or w.shape[1] != zs.size:
            raise ValueError('zs, chis and weights disagree in length')
        if keep is not None:
            dest = np.ascontiguousarray(keep[2], dtype=np.int32)
            if dest.shape != (w.shape[0],):
                raise ValueError('keep: one destination row per batch entry')
            _ffi.check(_ffi.lib().lbg_limber_keep(ctx.handle, w.shape[0], ells.size, _ffi._dp(ells), zs.size, _ffi._dp(chis), _ffi._dp(w),
                                                  prow.ctypes.data_as(_ffi.c_int_p), int(oob), int(keep[0]), int(keep[1]),
                                                  dest.ctypes.data_as(_ffi.c_int_p)))
            return None
        out = np.empty((w.shape[0], ells.size))
        _ffi.check(_ffi.lib().lbg_limber(ctx.handle, w.shape[0], ells.size, _ffi._dp(ells), zs.size, _ffi._dp(chis), _ffi._dp(w),
                                         prow.ctypes.data_as(_ffi.c_int_p), int(oob), _ffi._dp(out)))
        return out

    def limber_rows(self, ells, chis, weights, k, rows, prow, oob=1):
        """Limber sum over caller-supplied rows P_r(k) (modi/lensing.py `kp` tables); replaces the device table."""
        ctx = _ffi.context(self.device)
        k, rows = _ffi.f64(k), _ffi.f64(np.atleast_2d(rows))
        if rows.shape[1] != k.size:
            raise ValueError('k and P disagree in length')
        _ffi.check(_ffi.lib().lbg_pk_upload_rows(ctx.handle, k.size, _ffi._dp(k), rows.shape[0], _ffi._dp(rows)))
        _owner[ctx.device] = None                              # the linear table must be uploaded again before HALOFIT rows
        self._rows_key = None
        self._slots = {}
        ells, chis, w = _ffi.f64(np.atleast_1d(ells)), _ffi.f64(np.atleast_1d(chis)), _ffi.f64(np.atleast_2d(weights))
        prow = np.ascontiguousarray(prow, dtype=np.int32)
        if w.shape[1] != chis.size or prow.size != chis.size:
            raise ValueError('chis, weights and prow disagree in length')
        out = np.empty((w.shape[0], ells.size))
        _ffi.check(_ffi.lib().lbg_limber(ctx.handle, w.shape[0], ells.size, _ffi._dp(ells), chis.size, _ffi._dp(chis), _ffi._dp(w),
                                         prow.ctypes.data_as(_ffi.c_int_p), int(oob), _ffi._dp(out)))
        return out

    def pk_eval(self, z, k, nonlinear=True, oob=0):
        """np.interp(k, klin, P(z)) on the device: the `Pk_interps[...] (z, k)` callable of pmh.py:18-37."""
        self.build_rows([z], nonlinear=nonlinear)
        ctx = self._ctx()
        ka = _ffi.f64(np.atleast_1d(k))
        out = np.empty(ka.size)
        _ffi.check(_ffi.lib().lbg_pk_eval(ctx.handle, 0, ka.size, _ffi._dp(ka.ravel()), int(oob), _ffi._dp(out)))
        return out.reshape(ka.shape)
