"""CPU: the oracle restatements reproduce the golden vectors written by the UNMODIFIED reference
(oracle/make_golden.py, run in the build container).  This is what pins the oracle."""
import os

import numpy as np
import pytest

from oracle import halofit as ohf, limber as olim, n0 as on0
from oracle.shim.astropy.cosmology import FlatLambdaCDM


def _g(golden_dir, name):
    return np.load(os.path.join(golden_dir, name))


def test_n0_root_against_reference_vector(golden_dir):
    g = _g(golden_dir, 'n0_root_so_ttee.npz')     # 1 L: ~4 s on one core
    li, ni = on0.cl_interps(g['ell'], g['lensed']), on0.cl_interps(g['ell'], g['unlensed'])
    got = on0.nkk_root(li, ni, g['Ls'], tuple(str(t) for t in g['terms']), float(g['thetab']), float(g['DeltaT']),
                       bool(g['iterative']), Tcmb=float(g['Tcmb']), cos_phi=g['cos_phi'], sin_phi=g['sin_phi'])
    assert np.max(np.abs(got / g['nkk_reference'] - 1)) < 1e-13


def test_halofit_against_reference_vectors(golden_dir):
    g = _g(golden_dir, 'halofit.npz')
    cosmo = FlatLambdaCDM(H0=100., Om0=float(g['om_m']), Ob0=float(g['om_b']))
    k, p = g['k'], g['plin']
    for i in (0, 3):
        z, d = g['zs'][i], g['D'][i]
        delta = p * k ** 3 / (2. * np.pi ** 2) * d ** 2.
        assert np.allclose(ohf.get_spec(k, delta, float(g['sig8'])), g['spec'][i], rtol=1e-12)
        assert np.max(np.abs(ohf.halofit(k, delta, float(g['sig8']), z, cosmo, True) / g['delta_nl'][i] - 1)) < 1e-13
        assert np.max(np.abs(ohf.halofit(k, delta, float(g['sig8']), z, cosmo, False) / g['delta_nl_smith'][i] - 1)) < 1e-13


def test_limber_cheap_cases_against_reference_vectors(golden_dir):
    """zeff (thin-slice) spectra, shot noise and the lensing kernel: one halofit call each."""
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200.synthetic import PARAMS
    g = _g(golden_dir, 'limber_root.npz')
    eng = Cosmology(M=PARAMS['om_m'], H0=100., B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=g['k'], plin=g['plin'])
    Dg = lambda z: float(eng.Dgrow(z=z))
    assert np.allclose([Dg(z) for z in g['zchk']], g['Dchk'], rtol=1e-13)      # the growth input did not drift
    cosmo = FlatLambdaCDM(H0=100., Om0=PARAMS['om_m'], Ob0=PARAMS['om_b'])
    pk = ohf.ModiPk(g['k'], g['plin'], Dg, cosmo, PARAMS['sig_8']).interps()
    bg = olim.Background(PARAMS)
    assert np.allclose(bg.chi(g['zchk']), g['chichk'], rtol=1e-14)
    from oracle.simps121 import simps

    def pz(z, z0=2.96, s=0.24):     # Gaussian_pz.py:11-29
        zs = np.arange(0., 10., 0.01)
        return np.exp(-(z - z0) ** 2. / (2. * s ** 2.)) / simps(np.exp(-(zs - z0) ** 2. / (2. * s ** 2.)), dx=0.01)
    bz = lambda z: 1. + z
    L = g['Llls']
    zmin, zmax = float(g['zmin']), float(g['zmax'])
    assert np.max(np.abs(olim.Cgg(bg, pk, L, zmin, zmax, pz, bz, zeff=True) / g['cgg_zeff'] - 1)) < 1e-12
    assert np.max(np.abs(olim.Ckg(bg, pk, L, zmin, zmax, pz, bz, zeff=True) / g['ckg_zeff'] - 1)) < 1e-12
    assert np.max(np.abs(olim.Ngg(L, zmin, zmax, pz, 2181.0) / g['ngg'] - 1)) < 1e-13
    assert np.allclose(bg.lensing_kernel(np.array([0.5, 1.0, 3.0, 10.0])), g['lensing_kernel'], rtol=1e-14)


def test_cadence_fisher_against_reference_vector(golden_dir):
    """oracle/cadence.py vs the unmodified cadence.Fisher (print statements parenthesised in memory): both flavours."""
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200.synthetic import PARAMS
    from oracle import cadence as ocad
    from oracle.make_golden import cadence_inputs
    g = _g(golden_dir, 'cadence.npz')
    eng = Cosmology(M=PARAMS['om_m'], H0=100., B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=g['k'], plin=g['plin'])
    cosmo = FlatLambdaCDM(H0=100., Om0=PARAMS['om_m'], Ob0=PARAMS['om_b'])
    pk = ohf.ModiPk(g['k'], g['plin'], lambda z: float(eng.Dgrow(z=z)), cosmo, PARAMS['sig_8']).interps()
    bg = olim.Background(PARAMS)
    T = cadence_inputs()
    assert np.allclose([ocad.lenses_bz(bg, z) for z in g['zs']], g['lenses_bz'], rtol=1e-13)
    for name, sources in (('lenses', False), ('sources', True)):
        got = ocad.Fisher(bg, pk, g['L'].astype(np.float64), g['zs'], float(g['tNs']), float(g['tNp']), T['pz'], T['bz'],
                          fsky=float(g['fsky']), fover=float(g['fover']), sources=sources)
        assert np.max(np.abs(got / g['out_' + name] - 1)) < 1e-10, name


def test_eh_oracle_against_compiled_reference(golden_dir):
    """oracle/eh.py (float64) vs the reference's own EisensteinAndHu.c (float32): golden vectors everywhere, and the compiled
    library itself (oracle/_ref/libeh_ref.so, built by oracle/Makefile) wherever it is present."""
    from oracle import eh
    g = _g(golden_dir, 'eh.npz')
    for i, case in enumerate(g['cases']):
        args = tuple(case[:3]) + (int(case[3]),) + tuple(case[4:])
        cb, cbnu = eh.onek_hmpc(eh.set_cosm(*args), g['k'])
        assert np.max(np.abs(cb / g['tf_cb_%d' % i] - 1)) < 2e-6
        assert np.max(np.abs(cbnu / g['tf_cbnu_%d' % i] - 1)) < 2e-6
        if eh.ref_available():
            live = eh.ref_transfer(args, g['k'][::40])
            assert np.array_equal(live[0], g['tf_cb_%d' % i][::40])
    # SURVEY.md probe: T(k = 0.1 h/Mpc) = 0.1273156 for the RunPB cosmology at z = 0
    assert abs(eh.onek_hmpc(eh.set_cosm(0.292, 0.0468, 0.0, 1, 0.708, 0.69, 0.0), np.array([0.1]))[0][0] - 0.1273156) < 2e-7


@pytest.mark.reference
def test_reference_modules_import_under_shim():
    from oracle import refload
    if not refload.available():
        pytest.skip('/root/reference not present (GPU box)')
    for m in ('Nkk', 'Cgg', 'Ckg', 'lensing', 'halofit', 'nkk', 'snr', 'prep_Llls'):
        assert refload.load(m).__file__.startswith(refload.REF)


def test_fisher_oracle_against_reference_vectors(golden_dir):
    """snr / get_ClsCov / Fisher + bias loop / mqw.Fisher / rho_L restatements vs what the unmodified reference produced."""
    from oracle import fisher as ofi
    from oracle.make_golden import fisher_tracers
    g = _g(golden_dir, 'fisher.npz')
    z = np.zeros_like(g['C_kg'])
    Cls = {'gg': g['C_gg'], 'kg': g['C_kg'], 'gk': g['C_kg'], 'kk': g['C_kk']}
    Nls = {'gg': g['N_gg'], 'kg': z, 'gk': z, 'kk': g['N_kk']}
    xCls = {'gg': g['xC_gg'], 'kg': g['xC_kg'], 'gk': g['xC_kg'], 'kk': g['xC_kk']}
    xNls = {'gg': g['xN_gg'], 'kg': z, 'gk': z, 'kk': g['xN_kk']}
    L, nm, fsky = g['Llls'], g['nmodes'], float(g['fsky'])
    assert ofi.snr(L, g['C_kg'], g['vkg'], nm) == g['snr_all']
    assert ofi.snr(L, g['C_kg'], g['vkg'], nm, lmax=float(g['LllMax'])) == g['snr_cut']
    cov = ofi.get_ClsCov(L, Cls, Nls, fsky)
    for k, ref in zip(g['cov_keys'], g['cov']):
        assert np.array_equal(cov[str(k)], ref), k
    fb = ofi.fisher_bias(L, nm, Cls, Nls, xCls, xNls, fsky, float(g['fid_sig8']), float(g['fid_b1']), float(g['LllMax']))
    for key in ('Fisher', 'iFisher', 'interim', 'dtheta'):
        assert np.array_equal(fb[key], g[key]), key
    for i in range(g['yu_L'].size):
        assert ofi.rho_from_spectra(g['yu_ckk'][i], g['yu_kg'][i], g['yu_gg'][i], g['yu_ngg'][i]) == g['yu_rho'][i]
    # mqw.Fisher: per-shell scalars as mqw.py:24-41, then the oracle's matrix -> the reference's percent errors
    T = fisher_tracers()
    zs = np.arange(2.0 + 0.05, 4.0 + 0.05, 0.1)
    conv = lambda n: n / (np.pi / 180.) ** 2.
    bins = []
    for zee in zs:
        Ns, Np = conv(50.) * T['mqw_ps'](zee) * 0.1, conv(2000.) * T['mqw_pp'](zee) * 0.1
        bins.append([T['mqw_bp'](zee), T['mqw_bs'](zee), Np, Ns, Np, Ns, 0.1 * min(Np, Ns)])
    bins = np.array(bins, dtype=np.float64)
    F, Fii, S = ofi.mqw_fisher(g['mqw_cgg'], g['mqw_L'].astype(float), bins, 0.35)
    out = np.empty(40)
    out[0::2] = 100. * np.sqrt(1. / Fii) / bins[:, 2]
    out[1::2] = 100. * np.sqrt(np.diag(np.linalg.inv(F))) / bins[:, 2]
    assert np.max(np.abs(out / g['mqw_out'] - 1)) < 1e-12


def test_modi_n0_oracle_against_reference_vector(golden_dir):
    from scipy.interpolate import InterpolatedUnivariateSpline as IUS
    from lbgcmb_b200.synthetic import cmb_cl_tables
    g = _g(golden_dir, 'n0_modi.npz')
    ell, lens, nolens = cmb_cl_tables(5801)
    eu, el = ell[2:5801], ell[2:5301]
    thetab, delT = float(g['thetab']), float(g['delT'])
    n1 = on0.detector_noise(el, thetab, delT, 'TT', 2.786)
    lint = np.linspace(1., 3000., 10000)
    args = (IUS(eu, nolens[0, 2:5801]), IUS(eu, nolens[0, 2:5801]), IUS(el, lens[0, 2:5301] + n1), IUS(el, lens[0, 2:5301] + n1))
    got = on0.modi_N_fisher(g['Ls'][1], lint, *args, 'TT')          # ~2 s
    assert abs(got / g['int_TT'][1] - 1) < 1e-13


def test_modi_lensing_oracle_against_reference_vectors(golden_dir):
    """ModiCosmology + LensingX restatements vs the reference's own modi/cosmology.py + modi/lensing.py outputs
    (the z-resolved clz cases take 200 HALOFIT calls and are pinned by oracle/make_golden.py, which prints 0.0 for them)."""
    from oracle import modi_lensing as oml
    from lbgcmb_b200.synthetic import PARAMS
    g = _g(golden_dir, 'modi_lensing.npz')
    c = oml.ModiCosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=g['k'], plin=g['plin'])
    assert np.array_equal([c.Dgrow(z=z) for z in g['zchk']], g['Dchk'])
    assert np.array_equal([c.chia(z=z) for z in g['zchk']], g['chichk'])
    assert np.array_equal(c.pkanlin(z=1.0)[1], g['pknl_z1'])
    dndz = lambda z: z ** 2. * np.exp(-(z / 0.5) ** 1.0)
    dndz2 = lambda z: z ** 1.5 * np.exp(-(z / 0.4) ** 1.2)
    lx = oml.LensingX(z=1.0, dndz=dndz, cosmo=c, l=g['l'])
    hh = [g['k'], g['pknl_z1'] * 1.438 ** 2.]
    hm = [g['k'], g['pknl_z1'] * 1.438]
    assert np.array_equal(lx.cl_zeff(hh, auto=True), g['clggzeff'])
    assert np.array_equal(lx.cl_zeff(hm, auto=False), g['clkgzeff'])
    assert np.array_equal(lx.cl_zeff_bitracer(hh, dndz2), g['clbitracer'])
    assert np.array_equal(lx.cl_kkz(lambda z: [g['k'], c.pkanlin(z=z)[1]]), g['clkkz'])
    for name in ('kerg', 'kercmb', 'dggwtsdN', 'dkgwtsdN', 'chis'):
        assert np.array_equal(getattr(lx, name), g[name]), name


def test_romberg_restatements_agree():
    from oracle.romberg121 import romberg as r1
    from lbgcmb_b200.quadrature import romberg as r2
    f = lambda x: np.exp(-x * x) * np.cos(3 * x)
    for a, b in ((0., 1.), (-20., 0.), (0., 0.), (-2., 5.)):
        assert r1(f, a, b, vec_func=True) == r2(f, a, b)
    assert abs(r2(lambda x: x ** 3, 0., 2.) - 4.0) < 1e-13        # exact after one Richardson step
