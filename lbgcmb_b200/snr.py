"""Cumulative signal-to-noise; mirrors reference snr.py:21-28 (`snr`, `zeros`).  The masked sum over multipoles runs on
the device (csrc/fisher.cu snr_kernel) through lbg_snr."""
import numpy as np

from lbgcmb_b200 import _ffi


def snr(Llls, Cls, Vars, nmodes, lmax=1.e5, device=None):
    ctx = _ffi.context(device)
    L, C, V, N = (_ffi.f64(x) for x in (Llls, Cls, Vars, nmodes))
    if not (L.shape == C.shape == V.shape == N.shape) or L.ndim != 1:
        raise ValueError('Llls, Cls, Vars and nmodes must be 1-d arrays of one length')
    out = np.empty(1)
    _ffi.check(_ffi.lib().lbg_snr(ctx.handle, L.size, _ffi._dp(L), _ffi._dp(C), _ffi._dp(V), _ffi._dp(N), float(lmax), _ffi._dp(out)))
    return np.float64(out[0])


def zeros(arg):
    return np.zeros_like(arg)


def snr_grid(Llls, ckg, ckk, cgg, nkk, fsky, rho, ngg, nmodes, lmax=1.e5, device=None):
    """Cumulative S/N of C_kg for a whole grid of designs in one launch (lbg_snr_grid): experiment e (nkk[e], fsky[e]) x delensing
    curve r (rho[r], an array over Llls; zeros = none) x shot noise ngg[d].  out[e, r, d] equals
    snr(Llls, ckg, var_Ckg(..., nkk=nkk[e], ngg=ngg[d], fsky=fsky[e], rho=rho_r), nmodes, lmax) bit for bit (Ckg.py:61-87, snr.py:21-25)."""
    ctx = _ffi.context(device)
    L, kg, kk, gg, nm = (_ffi.f64(x) for x in (Llls, ckg, ckk, cgg, nmodes))
    nkk, rho = _ffi.f64(np.atleast_2d(nkk)), _ffi.f64(np.atleast_2d(rho))
    fsky, ngg = _ffi.f64(np.atleast_1d(fsky)), _ffi.f64(np.atleast_1d(ngg))
    n = L.size
    if L.ndim != 1 or not (kg.shape == kk.shape == gg.shape == nm.shape == (n,)) or nkk.shape[1] != n or rho.shape[1] != n \
            or fsky.shape != (nkk.shape[0],) or ngg.ndim != 1:
        raise ValueError('Llls, ckg, ckk, cgg, nmodes: one length; nkk [n_exp, n]; rho [n_rho, n]; fsky [n_exp]; ngg [n_dil]')
    out = np.empty((nkk.shape[0], rho.shape[0], ngg.size))
    _ffi.check(_ffi.lib().lbg_snr_grid(ctx.handle, n, _ffi._dp(L), _ffi._dp(nm), _ffi._dp(kk), _ffi._dp(gg), _ffi._dp(kg), nkk.shape[0],
                                       _ffi._dp(nkk), _ffi._dp(fsky), rho.shape[0], _ffi._dp(rho), ngg.size, _ffi._dp(ngg), float(lmax),
                                       _ffi._dp(out)))
    return out


def dilution_sweep(Pk_interps, lensCl_interps, nolensCl_interps, Llls, nmodes, zmin, zmax, pz, bz, nbar, experiments, rhos, dfactors,
                   Lmax, zeff=True, device=None):
    """The loop nest of the reference's snr.py driver (:107-133) as three device stages: for every CMB experiment
    (`experiments`: {name: {'fsky', 'thetab', 'DeltaT', 'iterative'}}, bolometers.py) x delensing efficiency (`rhos`: callables of L
    or None, snr.py:117 uses [zeros, rho]) x dilution factor (nbar -> dfactor * nbar, :108), the cumulative S/N of C_kg up to Lmax
    divided by sqrt(fsky) (:125).  The spectra are projected once, N0 once per experiment (levels of one beam share a pass of the
    multi-level kernel), and the 3 x 2 x 12 designs of the reference are ONE lbg_snr_grid launch.
    Returns (results [n_exp, n_rho, n_dil], names) with results[e, r, d] what the reference appends at :127."""
    from lbgcmb_b200.Cgg import Cgg, Ngg
    from lbgcmb_b200.Ckg import Ckg
    from lbgcmb_b200.lensing import Ckk
    from lbgcmb_b200 import Nkk as _Nkk
    Llls = np.asarray(Llls)
    names = list(experiments)
    ckg = Ckg(Pk_interps, Llls, zmin, zmax, pz, bz, zeff=zeff)
    ckk = Ckk(Pk_interps, Llls)
    cgg = Cgg(Pk_interps, Llls, zmin, zmax, pz, bz, zeff=zeff)
    nkk = np.empty((len(names), Llls.size))
    by_beam = {}
    for e, name in enumerate(names):
        by_beam.setdefault(float(experiments[name]['thetab']), []).append(e)
    for thetab, idx in by_beam.items():
        dts = [float(experiments[names[e]]['DeltaT']) for e in idx]
        its = [bool(experiments[names[e]]['iterative']) for e in idx]
        try:
            curves = _Nkk.n0_levels(lensCl_interps, nolensCl_interps, Llls, thetab, dts, its, device=device)[0] if len(idx) > 1 else None
        except ValueError:                                                  # a level the multi-level kernel does not serve: one call each
            curves = None
        for j, e in enumerate(idx):
            nkk[e] = curves[j] if curves is not None else _Nkk.Nkk(lensCl_interps, nolensCl_interps, Llls, ['TT', 'TE', 'EE', 'EB'], thetab,
                                                                   dts[j], iterative=its[j], pickleit=False)
    rho = np.array([np.zeros(Llls.size) if r is None else np.asarray(r(Llls), dtype=np.float64) for r in rhos])
    ngg = np.array([Ngg(Llls, zmin, zmax, pz, f * nbar)[0] for f in dfactors])
    fsky = np.array([float(experiments[name]['fsky']) for name in names])
    out = snr_grid(Llls, ckg, ckk, cgg, nkk, fsky, rho, ngg, nmodes, lmax=Lmax, device=device)
    return out / np.sqrt(fsky)[:, None, None], names
