"""GPU parity: HALOFIT table + Limber kernels (csrc/limber.cu) through the mirrors of modi/halofit.py, Cgg.py, Ckg.py,
cmb/lensing.py vs the reference's golden vectors (tests/golden/halofit.npz, limber_root.npz) and the numpy oracle.
Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_halofit_golden(ctx, golden_dir):
    from lbgcmb_b200.cosmo import FlatLambdaCDM
    from lbgcmb_b200.modi import halofit as hf
    g = _g(golden_dir, 'halofit.npz')
    cosmo = FlatLambdaCDM(H0=100., Om0=float(g['om_m']), Ob0=float(g['om_b']))
    k, p = g['k'], g['plin']
    for i, (z, d) in enumerate(zip(g['zs'], g['D'])):
        delta = p * k ** 3 / (2. * np.pi ** 2) * d ** 2.
        spec = hf._get_spec(k, delta, float(g['sig8']))
        assert np.allclose(spec, g['spec'][i], rtol=1e-10), (z, spec, g['spec'][i])
        got = hf.halofit(k, delta, float(g['sig8']), z, cosmo, takahashi=True)
        assert np.max(np.abs(got / g['delta_nl'][i] - 1)) < RTOL, z
        got = hf.halofit(k, delta, float(g['sig8']), z, cosmo, takahashi=False)
        assert np.max(np.abs(got / g['delta_nl_smith'][i] - 1)) < RTOL, z


def test_halofit_specs_per_redshift_golden(ctx, golden_dir):
    """Cosmology.halofit_specs (device: one inverse fit at z = 0 + a per-redshift FITPACK-order forward fit, csrc/halofit.cu) against
    the UNMODIFIED reference's `_get_spec` at every golden redshift.  k_nl and n_eff: 1e-12.  The curvature is the second derivative of
    an interpolating quintic spline on a 0.0092 grid: the fit runs in FITPACK's own operation order and sigma_0^2 is summed in the
    reference's own order (tests/test_fitpack_cpu.py: bit-identical on identical samples), so what is left is ONE rounding per sample
    -- the reference scales every term of the k sum by D^2, the library scales the sum -- times ~1e4: |dC| <= 3e-11 (2e-10 relative
    at z = 3 where C = 0.13; any libm's log moves the reference's own value by as much).  The non-linear power built from them --
    the quantity that enters C_l -- is pinned to the reference's `halofit` at 1e-10.  The table built in the same call equals the
    explicit-row route."""
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200.synthetic import PARAMS
    g = _g(golden_dir, 'halofit.npz')
    eng = Cosmology(M=PARAMS['om_m'], H0=100., B=PARAMS['om_b'], sig8=float(g['sig8']), klin=g['k'], plin=g['plin'])
    got = eng.halofit_specs(g['zs'], g['D'] ** 2.)
    assert got.shape == (g['zs'].size, 3)
    rel = np.abs(got / g['spec'] - 1)
    assert np.max(rel[:, :2]) < 1e-12, rel
    assert np.max(np.abs(got[:, 2] - g['spec'][:, 2])) < 3e-11 and np.max(rel[:, 2]) < 3e-10, rel
    one = np.array([eng.halofit_spec(z) for z in g['zs']])            # per-z host route (both fits per redshift, as the reference)
    assert np.max(np.abs(one[:, :2] / g['spec'][:, :2] - 1)) < 1e-12 and np.max(np.abs(one[:, 2] - g['spec'][:, 2])) < 3e-11
    k = g['k']
    rows_nl = eng.build_rows(g['zs'], want_host=True)                 # device route: specs + rows in one call
    delta_nl = rows_nl * k ** 3 / (2. * np.pi ** 2)
    assert np.max(np.abs(delta_nl / g['delta_nl'] - 1)) < RTOL
    # rows built on the device in the same call == rows from explicitly passed parameters
    zs = np.array([0.0, 0.7, 2.5, 4.0, 6.0])
    a = eng.build_rows(zs, want_host=True)
    D2 = np.array([float(eng.Dgrow(z=z)) ** 2. for z in zs])
    rows = np.zeros((zs.size, 8))
    rows[:, 0] = D2
    rows[:, 1] = 1.0
    hf = zs <= 4.0
    rows[hf, 1:4] = eng.halofit_specs(zs[hf], D2[hf])
    rows[hf, 4], rows[hf, 5], rows[hf, 6] = eng.cosmo.Om(zs[hf]), eng.cosmo.Ode(zs[hf]), 1.0
    from lbgcmb_b200 import _ffi
    b = np.empty_like(a)
    _ffi.check(_ffi.lib().lbg_pk_build_table(eng._ctx().handle, zs.size, _ffi._dp(_ffi.f64(rows)), 0.0, eng.cosmo.Om0, -1.0, 1, _ffi._dp(b)))
    eng.invalidate_device()
    assert np.max(np.abs(a / b - 1)) < 1e-13


def test_two_tables_in_a_row_do_not_share_device_state(ctx, golden_dir):
    """Two Cosmology / PkTable objects built one after the other in a helper (CPython hands the second the id() of the first):
    ownership of the device table is by a never-reused token + a digest of the table, so each object's spectra are its own; a
    table edited in place is uploaded again."""
    from lbgcmb_b200 import pmh
    g = _g(golden_dir, 'limber_root.npz')
    L = np.array([100., 1000.])

    def ckk_of(scale, n_k=None):
        from lbgcmb_b200.lensing import Ckk
        k, p = g['k'], g['plin'] * scale
        if n_k:
            k, p = k[::2][:n_k], p[::2][:n_k]
        return Ckk(pmh.get_PkInterps(pmh.PkTable(k, p)), L)
    a = ckk_of(1.0)
    b = ckk_of(1.21)                       # same k grid, different P: linear part scales by 1.21, the non-linear part by more
    assert np.all(b > 1.2 * a)
    c = ckk_of(1.0)
    assert np.array_equal(c, a)
    d = ckk_of(1.0, n_k=900)               # another k grid length right after: no stale table, no buffer mismatch
    assert np.all(np.isfinite(d)) and np.max(np.abs(d / a - 1)) < 0.05
    T = pmh.PkTable(g['k'], g['plin'].copy())
    from lbgcmb_b200.lensing import Ckk
    Pk = pmh.get_PkInterps(T)
    a1 = Ckk(Pk, L)
    T.engine.plin *= 1.21                  # edited in place
    assert np.array_equal(Ckk(Pk, L), b) and np.array_equal(a1, a)


def test_sigma2_power_law_closed_form(ctx):
    """sigma^2(R) for Delta^2 = k^n on a wide grid: 0.5 Gamma(n/2) R^-n (SURVEY.md 8(c) closed forms)."""
    from math import gamma
    from lbgcmb_b200.modi.halofit import _DeviceTable
    k = np.exp(np.linspace(np.log(1e-6), np.log(1e4), 4001))
    for n in (2.0, 3.5):      # n >= 2 keeps the k < 1e-6 truncation below 1e-12
        tab = _DeviceTable(k, k ** n)
        R = np.array([0.1, 1.0, 7.0])
        assert np.allclose(tab.sigma2(R), 0.5 * gamma(n / 2) * R ** -n, rtol=1e-9)


def _setup(g):
    from lbgcmb_b200 import pmh
    from lbgcmb_b200.Gaussian_pz import Gaussian_pz
    T = pmh.PkTable(g['k'], g['plin'])
    Pk = pmh.get_PkInterps(T)
    assert np.allclose([float(T.engine.Dgrow(z=z)) for z in g['zchk']], g['Dchk'], rtol=1e-13)
    pz_u = lambda z: Gaussian_pz(z, 2.96, 0.24)
    pz_g = lambda z: Gaussian_pz(z, 3.8, 0.3)
    return Pk, pz_u, pz_g, (lambda z: 1. + z)


def test_limber_root_golden(ctx, golden_dir):
    from lbgcmb_b200 import Cgg, Ckg, lensing
    g = _g(golden_dir, 'limber_root.npz')
    Pk, pz_u, pz_g, bz = _setup(g)
    L, zmin, zmax = g['Llls'], float(g['zmin']), float(g['zmax'])
    got = {
        'cgg_zeff': Cgg.Cgg(Pk, L, zmin, zmax, pz_u, bz, zeff=True),
        'cgg_full': Cgg.Cgg(Pk, L, zmin, zmax, pz_u, bz, zeff=False),
        'cgg_ug_full': Cgg.Cgg(Pk, L, 1.8, 5.8, pz_u, bz, survey_pz2=pz_g, bz2=bz, zeff=False),
        'ckg_zeff': Ckg.Ckg(Pk, L, zmin, zmax, pz_u, bz, zeff=True),
        'ckg_full': Ckg.Ckg(Pk, L, zmin, zmax, pz_u, bz, zeff=False),
        'ckk': lensing.Ckk(Pk, L),
        'ngg': Cgg.Ngg(L, zmin, zmax, pz_u, 2181.0),
        'lensing_kernel': lensing.lensing_kernel(np.array([0.5, 1.0, 3.0, 10.0])),
    }
    for name, v in got.items():
        rel = np.max(np.abs(v / g[name] - 1))
        assert rel < RTOL, '%s rel=%.3e' % (name, rel)


def test_pk_callable_matches_oracle_and_rejects_foreign(ctx, golden_dir):
    from lbgcmb_b200 import Cgg, pmh
    from lbgcmb_b200.cosmo import FlatLambdaCDM
    from lbgcmb_b200.synthetic import PARAMS
    from oracle import halofit as ohf
    g = _g(golden_dir, 'limber_root.npz')
    Pk, pz_u, _, bz = _setup(g)
    eng = Pk['camb_nlin'].engine
    cosmo = FlatLambdaCDM(H0=100., Om0=PARAMS['om_m'], Ob0=PARAMS['om_b'])
    ora = ohf.ModiPk(g['k'], g['plin'], lambda z: float(eng.Dgrow(z=z)), cosmo, PARAMS['sig_8']).interps()
    ks = np.concatenate([[1e-6, 1e3], np.exp(np.linspace(np.log(2e-5), np.log(90.), 500))])
    for z in (0.3, 2.96, 7.0):
        for name in ('camb_lin', 'camb_nlin'):
            a, b = Pk[name](z, ks), ora[name](z, ks)
            assert a[0] == 0.0 and a[1] == 0.0                         # zero outside the table (left = right = 0)
            assert np.max(np.abs(a[2:] / b[2:] - 1)) < RTOL, (z, name)
    assert isinstance(Pk['camb_nlin'](1.0, 0.1), float)
    with pytest.raises(ValueError):
        pmh.Pmm(Pk, ks, 1.0, type='weird')
    with pytest.raises(TypeError):
        Cgg.Cgg({'camb_nlin': lambda z, k: k, 'camb_lin': lambda z, k: k}, g['Llls'], 0.96, 4.96, pz_u, bz)


def test_limber_full_grid_properties(ctx, golden_dir):
    """4951 multipoles: agrees with the 60-l grid where they coincide; linear in the bias; batching is consistent."""
    from lbgcmb_b200 import Ckg
    g = _g(golden_dir, 'limber_root.npz')
    Pk, pz_u, _, bz = _setup(g)
    Lfull = np.arange(50, 5001)
    full = Ckg.Ckg(Pk, Lfull, 0.96, 4.96, pz_u, bz, zeff=False)
    sub = Ckg.Ckg(Pk, g['Llls'], 0.96, 4.96, pz_u, bz, zeff=False)
    assert np.array_equal(full[g['Llls'] - 50], sub)
    twice = Ckg.Ckg(Pk, g['Llls'], 0.96, 4.96, pz_u, lambda z: 2. * (1. + z), zeff=False)
    assert np.max(np.abs(twice / (2. * sub) - 1)) < 1e-14
    eng = Pk['camb_nlin'].engine
    zs = np.array([1.0, 2.0, 3.0])
    chis = np.array([eng.chia(z=z) for z in zs])
    w = np.random.default_rng(0).random((19, 3))
    batched = eng.limber(Lfull[:300], zs, chis, w)
    for b in (0, 7, 8, 18):
        assert np.array_equal(batched[b], eng.limber(Lfull[:300], zs, chis, w[b:b + 1])[0])
    assert eng.limber(np.array([]), zs, chis, w).shape == (19, 0)


def test_limber_kernel_random_tables_vs_numpy(ctx):
    """lbg_limber on caller-supplied rows (lbg_pk_upload_rows) against np.interp sums: non-uniform k grid, unsorted chi (the cell
    search falls back to bisection after every jump), nodes below / above / exactly on the ends of the k grid in both
    out-of-bounds modes, zero weights, batch sizes that exercise the single-entry and the 8-entry kernels with a ragged tail."""
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200.synthetic import eh_like_plin
    rng = np.random.default_rng(5)
    k0, p0 = eh_like_plin()
    eng = Cosmology(M=0.292, B=0.0468, klin=k0, plin=p0)
    for n_k, n_z, n_ell, n_b in ((37, 5, 3, 1), (500, 130, 257, 1), (1999, 700, 400, 3), (64, 1000, 129, 9), (2, 70, 50, 8)):
        k = np.sort(10. ** rng.uniform(-4, 1, n_k))
        k[1:] = np.maximum(k[1:], k[:-1] * (1 + 1e-12) + 1e-300)
        rows = 10. ** rng.uniform(-2, 4, (n_z, n_k))
        ell = np.sort(rng.uniform(1., 3000., n_ell))
        chi = 10. ** rng.uniform(0.5, 4.5, n_z)                              # unsorted: x jumps across the grid from node to node
        chi[0] = (ell[0] + 0.5) / k[-1]                                      # x close to / on the last knot for the first multipole
        chi[-1] = (ell[-1] + 0.5) / (0.5 * k[0])                              # x below the grid for every multipole
        w = rng.normal(size=(n_b, n_z)) * (rng.uniform(size=(n_b, n_z)) > 0.3)
        w[:, n_z // 2] = 0.0                                                 # a node no batch entry weights
        prow = rng.permutation(n_z).astype(np.int32)
        for oob in (0, 1):
            got = eng.limber_rows(ell, chi, w, k, rows, prow, oob=oob)
            x = (ell[None, :] + 0.5) / chi[:, None]
            want = np.zeros((n_b, n_ell))
            scale = np.zeros((n_b, n_ell))
            for iz in range(n_z):
                pv = np.interp(x[iz], k, rows[prow[iz]], left=None if oob else 0.0, right=None if oob else 0.0)
                want += w[:, iz, None] * pv[None, :]
                scale += np.abs(w[:, iz, None] * pv[None, :])
            assert got.shape == want.shape
            assert np.max(np.abs(got - want) / (scale + 1e-300)) < 1e-13, (n_k, n_z, oob)


def test_limber_mirrors_unusual_inputs_against_oracle(ctx):
    """Cgg / Ckg / Ckk mirrors against the numpy oracle on inputs the golden vectors do not hold: multipoles from 0 to 1e5 (k
    above the table: P = 0, prep_camb / np.interp fill), z ranges that give even and odd Simpson grids, a narrow and a wide
    slice, thin-slice (zeff) and full integrals, a cross-sample C_gg."""
    import bench
    from lbgcmb_b200 import Cgg, Ckg, lensing, pmh
    from lbgcmb_b200.synthetic import gaussian_pz, linz_bz
    k, p, _ = bench.forecast_inputs()
    of = bench.OracleForecast(k, p)
    ol = of.ol
    Pk = pmh.get_PkInterps(pmh.PkTable(k, p))
    L = np.array([0., 1., 2., 37., 500.5, 4999., 30000., 1e5])
    pz, pz2 = gaussian_pz(2.96, 0.24), gaussian_pz(3.4, 0.4)
    for zmin, zmax in ((2.0, 4.0), (2.005, 3.995), (0.01, 10.0), (2.9, 3.0)):
        for zeff in (True, False):
            a = Cgg.Cgg(Pk, L, zmin, zmax, pz, linz_bz, zeff=zeff)
            b = ol.Cgg(of.bg, of.Pk, L, zmin, zmax, pz, linz_bz, zeff=zeff)
            assert np.allclose(a, b, rtol=RTOL, atol=0), ('cgg', zmin, zmax, zeff, np.max(np.abs(a - b) / np.abs(b).max()))
            a = Ckg.Ckg(Pk, L, zmin, zmax, pz, linz_bz, zeff=zeff)
            b = ol.Ckg(of.bg, of.Pk, L, zmin, zmax, pz, linz_bz, zeff=zeff)
            assert np.allclose(a, b, rtol=RTOL, atol=0), ('ckg', zmin, zmax, zeff)
    a = Cgg.Cgg(Pk, L, 2.0, 5.0, pz, linz_bz, survey_pz2=pz2, bz2=linz_bz, zeff=False)
    b = ol.Cgg(of.bg, of.Pk, L, 2.0, 5.0, pz, linz_bz, survey_pz2=pz2, bz2=linz_bz, zeff=False)
    assert np.allclose(a, b, rtol=RTOL, atol=0)
    assert np.allclose(lensing.Ckk(Pk, L), ol.Ckk(of.bg, of.Pk, L), rtol=RTOL, atol=0)
