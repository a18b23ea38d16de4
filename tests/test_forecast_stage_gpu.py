"""GPU parity of bench.py's forecast stage (Limber C_kk / C_kg / C_gg with HALOFIT rows on ~1000 / 999 / 2000 z nodes,
Gaussian covariance, cumulative S/N, (sigma8, b1) Fisher matrix and interloper bias for the four dropout bands) against the
numpy oracle behind the same call surface (bench.OracleForecast: oracle/limber.py, oracle/halofit.py, oracle/fisher.py).

Tolerance: relative 1e-10 on every spectrum; the S/N and Fisher entries are sums of those and are held to 1e-9, the bias
vector (a difference of nearly equal spectra pushed through a 2x2 inverse) to 1e-7 of its own scale.
This is also the wide check of the shared-spline HALOFIT glue (Cosmology.halofit_specs): ~310 redshifts below z = 4."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope='module')
def setup():
    import bench
    from lbgcmb_b200 import pmh
    kk, pp, bands = bench.forecast_inputs()
    Ls = np.array([50., 51., 120., 333., 700., 1199., 1200., 2048., 3000., 5000.])
    nmodes = np.arange(1., Ls.size + 1.)
    nkk = 1e-8 * (1. + Ls / 1000.) ** 2
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp))
    of = bench.OracleForecast(kk, pp)
    return bench, bands, Ls, nmodes, nkk, Pk, of


def test_spectra_per_band(ctx, setup):
    bench, bands, Ls, nmodes, nkk, Pk, of = setup
    from lbgcmb_b200 import interloper
    for b in bands:
        for pz, nbar in ((b['pz'], b['nbar']), (b['pzz'], b['nbar_wint'])):
            got, gotN = interloper.get_allCls(Pk, Ls, nbar, bench.FSKY, 0.01, 10.0, pz, b['bz'], zeff=False, nkk=nkk)
            ref, refN = of.get_allCls(of.Pk, Ls, nbar, bench.FSKY, 0.01, 10.0, pz, b['bz'], zeff=False, nkk=nkk)
            for key in ('gg', 'kg', 'kk'):
                rel = np.max(np.abs(got[key] / ref[key] - 1))
                assert rel < 1e-10, (b['name'], key, rel)
            assert np.max(np.abs(gotN['gg'] / refN['gg'] - 1)) < 1e-13


def test_stage_rows(ctx, setup):
    bench, bands, Ls, nmodes, nkk, Pk, of = setup
    from lbgcmb_b200 import interloper, snr
    got = bench.forecast_stage((interloper, snr), Pk, bands, Ls, nmodes, nkk)
    ref = bench.forecast_stage((of, of), of.Pk, bands, Ls, nmodes, nkk)
    assert got.shape == ref.shape == (len(bands), 7)
    assert np.max(np.abs(got[:, :5] / ref[:, :5] - 1)) < 1e-9          # S/N and the four Fisher entries
    scale = np.max(np.abs(ref[:, 5:]), axis=1, keepdims=True)
    assert np.max(np.abs(got[:, 5:] - ref[:, 5:]) / scale) < 1e-7      # interloper bias (dsigma8, db1)


def test_kernel_totals_accumulate(ctx, setup):
    bench, bands, Ls, nmodes, nkk, Pk, of = setup
    from lbgcmb_b200 import lensing
    ctx.kernel_totals(reset=True)
    assert ctx.kernel_totals() == (0.0, 0)
    lensing.Ckk(Pk, Ls)
    ms1, n1 = ctx.kernel_totals()
    lensing.Ckk(Pk, Ls)
    ms2, n2 = ctx.kernel_totals()
    assert n1 >= 2 and n2 >= n1 + 2 and ms2 > ms1 > 0.0                # limber + reduce launches per call at least
