"""GPU tests of the tomographic cosmology-parameter Fisher driver (lbgcmb_b200/tomography.py; BASELINE configs[3]).
The reference has no such driver, so it is pinned piecewise: the per-bin Limber sums against the reference-pinned
LensingX mirror (modi/lensing.py:69-179), the finite-difference sigma8 derivative against the exact 2 C / sigma8 of linear theory
(interloper.py:65-74), the Fisher sum against the dense numpy restatement."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu

dndz = lambda z: z ** 2. * np.exp(-(z / 0.5) ** 1.0)
bz = lambda z: 1.1 + 0.4 * z
ELL = np.unique(np.round(np.logspace(np.log10(50.), np.log10(3000.), 40)))


def test_one_bin_equals_lensingx(ctx):
    from lbgcmb_b200 import tomography as tm
    from lbgcmb_b200.modi.lensing import LensingX, PkZ
    c = tm.make_cosmology(tm.FIDUCIAL)
    ckk, kg, gg = tm.field_spectra(c, ELL, [{'edges': [0.75, 1.25], 'dndz': dndz, 'bz': bz}], nodes_per_bin=99)
    lx = LensingX(z=1.0, dndz=dndz, cosmo=c, l=ELL, dz=0.5, nz=100)
    assert np.max(np.abs(kg[0] / lx.clz(PkZ(c, bias=bz, power=1), auto=False) - 1)) < 1e-12
    assert np.max(np.abs(gg[0] / lx.clz(PkZ(c, bias=bz, power=2), auto=True) - 1)) < 1e-12
    assert np.array_equal(ckk, lx.cl_kkz(PkZ(c)))
    assert np.all(ckk > 0) and np.all(kg > 0) and np.all(gg > 0)


def _tracers():
    return [{'name': 'lo', 'edges': np.linspace(0.4, 1.0, 4), 'dndz': dndz, 'bz': bz, 'nbar': 3e6},
            {'name': 'hi', 'edges': np.linspace(1.0, 2.0, 3), 'dndz': lambda z: np.exp(-0.5 * ((z - 1.5) / 0.4) ** 2.),
             'bz': lambda z: 2.0 + 0. * z, 'nbar': [4e6, 2e6]}]


def test_sigma8_derivative_and_fisher_sum(ctx):
    from lbgcmb_b200 import tomography as tm
    from oracle.fisher import field_fisher as dense
    nm = np.ones_like(ELL)
    nkk = 2e-8 * np.ones_like(ELL)
    r = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('sig8',), nonlinear=False)
    assert r['params'] == ['sig8', 'b_lo', 'b_hi'] and r['fields'] == 6 and r['n_cosmologies'] == 3
    s = r['spectra']
    nf = 6
    spectra = [(0, 0)] + [(0, i) for i in range(1, nf)] + [(i, i) for i in range(1, nf)]
    tot = np.concatenate([(s['kk'] + nkk)[None], s['kg'], s['gg'] + s['ngg'][:, None]])
    C = np.concatenate([s['kk'][None], s['kg'], s['gg']])
    dC = np.zeros((3, len(spectra), ELL.size))
    dC[0] = 2. * C / tm.FIDUCIAL['sig8']                          # linear theory: every spectrum scales as sigma8^2
    dC[1, 1:4], dC[1, 6:9] = s['kg'][:3], 2. * s['gg'][:3]
    dC[2, 4:6], dC[2, 9:11] = s['kg'][3:], 2. * s['gg'][3:]
    F0, _ = dense(nf, spectra, tot, dC, nm * (2. * ELL + 1.) * 0.4)
    assert np.max(np.abs(r['F'] - F0) / np.sqrt(np.outer(np.diag(F0), np.diag(F0)))) < 1e-9
    assert np.all(np.linalg.eigvalsh(r['F']) > 0) and np.allclose(r['cov'] @ r['F'], np.eye(3), atol=1e-8)
    assert np.array_equal(s['ngg'], 1. / np.array([3e6, 3e6, 3e6, 4e6, 2e6]))


def test_neutrino_mass_forecast(ctx):
    from lbgcmb_b200 import tomography as tm
    nm = np.ones_like(ELL)
    nkk = 2e-8 * np.ones_like(ELL)
    r = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('sig8', 'mnu', 'M'))
    assert r['params'] == ['sig8', 'mnu', 'M', 'b_lo', 'b_hi'] and r['n_cosmologies'] == 7
    F = r['F']
    assert np.all(np.isfinite(F)) and np.array_equal(F, F.T) and np.all(np.linalg.eigvalsh(F) > 0)
    assert np.all(r['sigma'] > 0) and np.all(r['sigma'] >= 1. / np.sqrt(np.diag(F)) * (1 - 1e-12))     # marginalised >= conditional
    # the same call with the multipoles "sharded" over one rank, and halved noise -> more information on every parameter
    r1 = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('sig8', 'mnu', 'M'), sharded=True)
    assert np.array_equal(r1['F'], F)
    r2 = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, 0.5 * nkk, cosmo_params=('sig8', 'mnu', 'M'))
    assert np.all(np.diag(r2['F']) > np.diag(F))
    with pytest.raises(ValueError):
        tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('w0',))
    with pytest.raises(ValueError):
        tm.tomographic_fisher(_tracers() * 4, ELL, nm, 0.4, nkk, cosmo_params=('sig8',))


def test_geometry_cache_is_per_background(ctx):
    """Distances / weight vectors are shared by cosmologies with the same (M, L, H0) and rebuilt for another background;
    cached and uncached calls return identical spectra."""
    from lbgcmb_b200 import tomography as tm
    tr = _tracers()
    fid = tm.FIDUCIAL
    geo = {}
    a = tm.field_spectra(tm.make_cosmology(fid), ELL, tr, geometry=geo)
    assert len(geo) == 1
    b = tm.field_spectra(tm.make_cosmology(dict(fid, sig8=0.83)), ELL, tr, geometry=geo)
    assert len(geo) == 1 and np.all(b[1] > a[1])                                   # same background, more power
    c = tm.field_spectra(tm.make_cosmology(dict(fid, M=0.30)), ELL, tr, geometry=geo)
    assert len(geo) == 2 and not np.array_equal(c[1], a[1])
    a2 = tm.field_spectra(tm.make_cosmology(fid), ELL, tr, geometry=geo)            # served from the cache
    a3 = tm.field_spectra(tm.make_cosmology(fid), ELL, tr, kk=False)                # no cache, no C_kk
    assert a3[0] is None
    for x, y, z in zip(a[1:], a2[1:], a3[1:]):
        assert np.array_equal(x, y) and np.array_equal(x, z)
    assert np.array_equal(a[0], a2[0])


def test_device_resident_route_equals_host_array_route(ctx):
    """Spectra kept on the device (lbg_limber_keep -> lbg_field_fisher_sets) against the same kernels fed through host arrays
    (lbg_limber -> numpy sums and differences -> lbg_field_fisher_amp): the same doubles, so F and the fiducial spectra agree bit for bit."""
    from lbgcmb_b200 import tomography as tm
    nm = np.linspace(1., 2., ELL.size)
    nkk = 2e-8 * (1. + ELL / 1000.)
    kw = dict(cosmo_params=('sig8', 'mnu'))
    dev = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, **kw)
    host = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, host_arrays=True, **kw)
    assert np.array_equal(dev['F'], host['F'])
    for key in ('kk', 'kg', 'gg'):
        assert np.array_equal(dev['spectra'][key], host['spectra'][key])
    lean = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, return_spectra=False, **kw)
    assert np.array_equal(lean['F'], dev['F']) and lean['spectra']['kg'] is None


def test_spectra_sets_entry_points(ctx):
    """lbg_spectra_set_fill / _get round trip, lbg_field_fisher_sets == lbg_field_fisher_amp on the same numbers, and the state errors."""
    from lbgcmb_b200 import fisher
    rng = np.random.default_rng(5)
    nf, n_ell = 4, 23
    spectra = [(0, 0)] + [(0, i) for i in range(1, nf)] + [(i, i) for i in range(1, nf)]
    ns = len(spectra)
    fid = np.abs(rng.normal(1., 0.1, (ns, n_ell))) * np.array([1.] + [0.2] * (nf - 1) + [1.] * (nf - 1))[:, None]
    up, dn = fid * rng.uniform(1.0, 1.1, fid.shape), fid * rng.uniform(0.9, 1.0, fid.shape)
    noise_c = np.concatenate([np.zeros(nf), rng.uniform(0.1, 0.3, nf - 1)])
    nkk = rng.uniform(0.1, 0.2, n_ell)
    for s, m in ((3, fid), (4, up), (5, dn)):                                   # a set filled row by row from curves
        fisher.spectra_set_fill(s, np.zeros(ns), n_ell, curves={r: m[r] for r in range(ns)})
        assert np.array_equal(fisher.spectra_set_get(s, 0, ns, n_ell), m)
    fisher.spectra_set_fill(6, noise_c, n_ell, curves={0: nkk})
    noise = np.repeat(noise_c[:, None], n_ell, axis=1)
    noise[0] = nkk
    assert np.array_equal(fisher.spectra_set_get(6, 0, ns, n_ell), noise)
    assert np.array_equal(fisher.spectra_set_get(6, 2, 3, n_ell), noise[2:5])
    amp = np.zeros((1, ns))
    amp[0, 1:nf], amp[0, nf:] = 1., 2.
    w = rng.uniform(1., 5., n_ell)
    h = 0.01
    got = fisher.field_fisher_sets(nf, spectra, n_ell, 3, 6, [(4, 5, 2. * h)], amp, w)
    dC = ((up - dn) / (2. * h))[None]
    want = fisher.field_fisher_amp(nf, spectra, fid + noise, dC, amp, fid, w)
    assert np.array_equal(got, want)
    none = fisher.field_fisher_sets(nf, spectra, n_ell, 3, None, [(4, 5, 2. * h)], amp, w)
    assert np.array_equal(none, fisher.field_fisher_amp(nf, spectra, fid, dC, amp, fid, w))
    with pytest.raises(Exception, match='holds'):                               # a set of another shape
        fisher.field_fisher_sets(nf, spectra, n_ell - 1, 3, 6, [(4, 5, 2. * h)], amp, w[:-1])
    with pytest.raises(ValueError):
        fisher.field_fisher_sets(nf, spectra, n_ell, 3, 25, [(4, 5, 2. * h)], amp, w)
    with pytest.raises(ValueError):
        fisher.field_fisher_sets(nf, spectra, n_ell, 3, 6, [(4, 5, 0.0)], amp, w)
    with pytest.raises(ValueError):
        fisher.spectra_set_get(6, ns - 1, 2, n_ell)


def test_fsigma8_derived_errors(ctx):
    """f sigma8(z) errors propagated from the (sig8, M) block of the forecast: positive, finite, and with device growth tables equal to
    the host-table route; sigma8-only forecasts give sigma(f sigma8) / f sigma8 = sigma(sig8) / sig8."""
    from lbgcmb_b200 import tomography as tm
    nm = np.ones_like(ELL)
    nkk = 2e-8 * np.ones_like(ELL)
    zs = np.array([0.5, 1.0, 2.0, 3.0])
    r = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('sig8', 'M'))
    d = tm.fsigma8_forecast(r, zs)
    assert d['sigma'].shape == (4,) and np.all(d['sigma'] > 0) and np.all(np.isfinite(d['sigma']))
    assert np.all(d['jacobian'][:, 2:] == 0.0) and np.all(d['jacobian'][:, 0] > 0)
    h = tm.fsigma8_forecast(r, zs[:2], growth='host')
    assert np.max(np.abs(h['fsigma8'] / d['fsigma8'][:2] - 1)) < 1e-9 and np.max(np.abs(h['sigma'] / d['sigma'][:2] - 1)) < 1e-6
    r1 = tm.tomographic_fisher(_tracers(), ELL, nm, 0.4, nkk, cosmo_params=('sig8',))
    d1 = tm.fsigma8_forecast(r1, zs)
    assert np.max(np.abs(d1['sigma'] / d1['fsigma8'] - r1['sigma'][0] / tm.FIDUCIAL['sig8'])) < 1e-10
