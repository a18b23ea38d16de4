"""GPU parity of the survey-design sweep (lbgcmb_b200/sweep.py, csrc/sweep.cu): the batched launches must reproduce the
per-design mirror calls (Cgg.Cgg / Ckg.Ckg / snr.snr / interloper.fisher_bias) and, through them, the numpy oracle."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def setup(cl_tables):
    import bench
    from lbgcmb_b200 import pmh
    from lbgcmb_b200.prep_Llls import prep_Llls
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    kk, pp, bands = bench.forecast_inputs()
    _, Ls, nmodes = prep_Llls(NLlls=24, Lmin=50., Lmax=5000., log10=True)
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp))
    samples = [{'pz': b['pz'], 'bz': b['bz'], 'nbar': b['nbar'], 'fid_b1': b['fid_b1'], 'LllMax': b['LllMax']} for b in bands[:3]]
    xsamples = [{'pz': b['pzz'], 'bz': b['bz'], 'nbar': b['nbar_wint']} for b in bands[:3]]
    li, ni = cl_interps_from_tables(*cl_tables)
    return bench, Pk, Ls.astype(np.float64), nmodes.astype(np.float64), samples, xsamples, li, ni


def test_limber_grid_matches_per_sample_calls(ctx, setup):
    from lbgcmb_b200 import Cgg, Ckg, sweep
    bench, Pk, Ls, nmodes, samples, xsamples, li, ni = setup
    cgg, ckg, ngg = sweep.limber_grid(Pk, Ls, samples)
    # the batched call builds the Limber weights on the device (block-tree sum in the p(z) renormalisation), the mirrors on the host
    # (numpy pairwise sum): same arithmetic, last-bit differences
    for i, s in enumerate(samples):
        assert np.max(np.abs(cgg[i] / Cgg.Cgg(Pk, Ls, 0.01, 10.0, s['pz'], s['bz'], zeff=False) - 1)) < 1e-13
        assert np.max(np.abs(ckg[i] / Ckg.Ckg(Pk, Ls, 0.01, 10.0, s['pz'], s['bz'], zeff=False) - 1)) < 1e-13
        assert ngg[i] == Cgg.Ngg(Ls, 0.01, 10.0, s['pz'], s['nbar'])[0]
    ckk, cgg2, ckg2, ngg2 = sweep.forecast_spectra(Pk, Ls, samples)
    from lbgcmb_b200 import lensing
    assert np.array_equal(ckk, lensing.Ckk(Pk, Ls)) and np.array_equal(cgg2, cgg) and np.array_equal(ckg2, ckg)


def test_forecast_grid_matches_mirrors(ctx, setup):
    from lbgcmb_b200 import interloper, snr, sweep, lensing
    bench, Pk, Ls, nmodes, samples, xsamples, li, ni = setup
    ckk = lensing.Ckk(Pk, Ls)
    cgg, ckg, ngg = sweep.limber_grid(Pk, Ls, samples)
    xcgg, xckg, xngg = sweep.limber_grid(Pk, Ls, xsamples)
    nkk = np.stack([1e-8 * (1. + Ls / 900.) ** 2, 3e-8 * (1. + Ls / 400.) ** 2])
    out = sweep.forecast_grid(Ls, nmodes, ckk, cgg, ckg, ngg, nkk, bench.FSKY, 0.819, [s['fid_b1'] for s in samples],
                              [s['LllMax'] for s in samples], xcgg=xcgg, xckg=xckg, xngg=xngg)
    assert out.shape == (2, 3, 8)
    z = np.zeros_like(Ls)
    for n in range(2):
        for g, s in enumerate(samples):
            Cls = {'gg': cgg[g], 'kg': ckg[g], 'gk': ckg[g], 'kk': ckk}
            Nls = {'gg': ngg[g] * np.ones_like(Ls), 'kg': z, 'gk': z, 'kk': nkk[n]}
            xCls = {'gg': xcgg[g], 'kg': xckg[g], 'gk': xckg[g], 'kk': ckk}
            xNls = {'gg': xngg[g] * np.ones_like(Ls), 'kg': z, 'gk': z, 'kk': nkk[n]}
            vkg = ((ckk + nkk[n]) * (cgg[g] + ngg[g]) + ckg[g] ** 2.) / ((2. * Ls + 1.) * bench.FSKY)
            sn = snr.snr(Ls, ckg[g], vkg, nmodes, lmax=s['LllMax'])
            fb = interloper.fisher_bias(Ls, nmodes, Cls, Nls, xCls, xNls, bench.FSKY, 0.819, s['fid_b1'], s['LllMax'], fused=False)
            ref = np.concatenate([[sn], fb['Fisher'].ravel(), fb['interim']])
            assert np.max(np.abs(out[n, g, :7] / ref - 1)) < 1e-12, (n, g, out[n, g], ref)
    nob = sweep.forecast_grid(Ls, nmodes, ckk, cgg, ckg, ngg, nkk, bench.FSKY, 0.819, [s['fid_b1'] for s in samples],
                              [s['LllMax'] for s in samples])
    assert np.array_equal(nob[:, :, :5], out[:, :, :5]) and np.all(nob[:, :, 5:] == 0.0)
    with pytest.raises(ValueError):
        sweep.forecast_grid(Ls, nmodes, ckk, cgg, ckg[:, :-1], ngg, nkk, bench.FSKY, 0.819, 1., 1000.)


def test_survey_sweep_end_to_end_vs_oracle(ctx, setup):
    """2 noise levels x 3 n(z): N0 on the device, everything else too; the oracle runs the same designs one by one."""
    from lbgcmb_b200 import sweep
    from oracle import n0 as on0
    bench, Pk, Ls, nmodes, samples, xsamples, li, ni = setup
    sel = np.array([0, 9, 17, 23])
    L4, nm4 = Ls[sel], nmodes[sel]
    noise = [(1.4, 1.0, True), (3.0, 6.0, False)]
    res = sweep.survey_sweep(li, ni, Pk, L4, nm4, noise, samples, bench.FSKY, xsamples=xsamples)
    assert res['out'].shape == (2, 3, 8) and res['nkk'].shape == (2, 4)
    ell, lens, nolens = bench.make_inputs()
    kk, pp, _ = bench.forecast_inputs()
    of = bench.OracleForecast(kk, pp)
    for n, cfg in enumerate(noise):
        nkk = on0.nkk_root(on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens), L4.astype(int), ('TT', 'TE', 'EE', 'EB'), *cfg)
        assert np.max(np.abs(res['nkk'][n] / nkk - 1)) < 1e-10
        for g, (s, x) in enumerate(zip(samples, xsamples)):
            Cls, Nls = of.get_allCls(of.Pk, L4, s['nbar'], bench.FSKY, 0.01, 10.0, s['pz'], s['bz'], nkk=nkk)
            xCls, xNls = of.get_allCls(of.Pk, L4, x['nbar'], bench.FSKY, 0.01, 10.0, x['pz'], x['bz'], nkk=nkk)
            vkg = ((Cls['kk'] + nkk) * (Cls['gg'] + Nls['gg']) + Cls['kg'] ** 2.) / ((2. * L4 + 1.) * bench.FSKY)
            sn = of.snr(L4, Cls['kg'], vkg, nm4, lmax=s['LllMax'])
            fb = of.fisher_bias(L4, nm4, Cls, Nls, xCls, xNls, bench.FSKY, 0.819, s['fid_b1'], s['LllMax'])
            ref = np.concatenate([[sn], fb['Fisher'].ravel()])
            assert np.max(np.abs(res['out'][n, g, :5] / ref - 1)) < 1e-9, (n, g)


def test_sharded_forecast_single_rank_matches_mirrors(ctx):
    """sweep.sharded_forecast on one rank (the whole L grid is its shard) = snr.snr + interloper.fisher_bias on the same inputs."""
    import bench
    from lbgcmb_b200 import Nkk, interloper, pmh, snr as snr_mod, sweep
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    li, ni = cl_interps_from_tables(*bench.make_inputs())
    kk, pp, bands = bench.forecast_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp))
    Ls = np.arange(50, 2400, 7)
    Lf = Ls.astype(np.float64)
    nm = np.full(Lf.size, 7.0)
    b = bands[1]
    smp = {'pz': b['pz'], 'bz': b['bz'], 'nbar': b['nbar'], 'fid_b1': b['fid_b1'], 'LllMax': b['LllMax']}
    xsmp = {'pz': b['pzz'], 'bz': b['bz'], 'nbar': b['nbar_wint']}
    got = sweep.sharded_forecast(li, ni, Pk, Ls, nm, (1.4, 7.0, False), smp, bench.FSKY, xsample=xsmp)
    nkk = Nkk.Nkk(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 7.0, iterative=False, pickleit=False)
    Cls, Nls = interloper.get_allCls(Pk, Lf, b['nbar'], bench.FSKY, 0.01, 10.0, b['pz'], b['bz'], zeff=False, nkk=nkk)
    xCls, xNls = interloper.get_allCls(Pk, Lf, b['nbar_wint'], bench.FSKY, 0.01, 10.0, b['pzz'], b['bz'], zeff=False, nkk=nkk)
    vkg = ((Cls['kk'] + nkk) * (Cls['gg'] + Nls['gg']) + Cls['kg'] ** 2.) / ((2. * Lf + 1.) * bench.FSKY)
    sn = snr_mod.snr(Lf, Cls['kg'], vkg, nm, lmax=b['LllMax'])
    fb = interloper.fisher_bias(Lf, nm, Cls, Nls, xCls, xNls, bench.FSKY, 0.819, b['fid_b1'], b['LllMax'], fused=False)
    assert np.array_equal(got['nkk'], nkk)
    assert np.max(np.abs(got['cgg'] / Cls['gg'] - 1)) < 1e-13 and np.max(np.abs(got['ckg'] / Cls['kg'] - 1)) < 1e-13
    assert abs(got['snr'] / sn - 1) < 1e-12
    assert np.max(np.abs(got['Fisher'] / fb['Fisher'] - 1)) < 1e-11
    assert np.max(np.abs(got['dtheta'] / fb['dtheta'] - 1)) < 1e-9
