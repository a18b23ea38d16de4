"""GPU parity: S/N, Gaussian covariance, Fisher / bias accumulation, N_p(z) Fisher and multi-tracer rho_L
(csrc/fisher.cu through the mirrors of snr.py, interloper.py, mqw/py/mqw.py, yu.py) vs golden vectors written by the
UNMODIFIED reference functions (tests/golden/fisher.npz, oracle/make_golden.py) and the numpy oracle (oracle/fisher.py).
Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


@pytest.fixture(scope='module')
def g(golden_dir):
    return np.load(os.path.join(golden_dir, 'fisher.npz'))


def _dicts(g):
    z = np.zeros_like(g['C_kg'])
    Cls = {'gg': g['C_gg'], 'kg': g['C_kg'], 'gk': g['C_kg'], 'kk': g['C_kk']}
    Nls = {'gg': g['N_gg'], 'kg': z, 'gk': z, 'kk': g['N_kk']}
    xCls = {'gg': g['xC_gg'], 'kg': g['xC_kg'], 'gk': g['xC_kg'], 'kk': g['xC_kk']}
    xNls = {'gg': g['xN_gg'], 'kg': z, 'gk': z, 'kk': g['xN_kk']}
    return Cls, Nls, xCls, xNls


def test_snr_golden_and_edges(ctx, g):
    from lbgcmb_b200.snr import snr, zeros
    L, nm = g['Llls'], g['nmodes']
    assert abs(snr(L, g['C_kg'], g['vkg'], nm) / g['snr_all'] - 1) < RTOL
    assert abs(snr(L, g['C_kg'], g['vkg'], nm, lmax=float(g['LllMax'])) / g['snr_cut'] - 1) < RTOL
    assert snr(L, g['C_kg'], g['vkg'], nm, lmax=10.) == 0.0               # nothing below the cut
    assert snr(np.array([]), np.array([]), np.array([]), np.array([])) == 0.0
    assert zeros(L).shape == L.shape
    with pytest.raises(ValueError):
        snr(L, g['C_kg'][:-1], g['vkg'], nm)


def test_get_ClsCov_bit_exact(ctx, g):
    """Elementwise products / sums / two divisions in the reference's order: identical bits."""
    from lbgcmb_b200.interloper import get_ClsCov
    Cls, Nls, _, _ = _dicts(g)
    cov = get_ClsCov(g['Llls'], Cls, Nls, float(g['fsky']))
    assert sorted(cov) == sorted(str(k) for k in g['cov_keys'])
    for k, ref in zip(g['cov_keys'], g['cov']):
        assert np.array_equal(cov[str(k)], ref), k


def test_fisher_bias_golden(ctx, g):
    from lbgcmb_b200.interloper import fisher_bias, dp_DL
    from oracle import fisher as ofi
    Cls, Nls, xCls, xNls = _dicts(g)
    args = (g['Llls'], g['nmodes'], Cls, Nls, xCls, xNls, float(g['fsky']), float(g['fid_sig8']), float(g['fid_b1']), float(g['LllMax']))
    got = fisher_bias(*args)                       # one launch (csrc/sweep.cu) when the inputs have get_allCls's shape
    gen = fisher_bias(*args, fused=False)          # get_ClsCov + fisher_accum
    for key in ('Fisher', 'iFisher', 'interim', 'dtheta'):
        assert np.max(np.abs(got[key] / g[key] - 1)) < RTOL, key
        assert np.max(np.abs(gen[key] / g[key] - 1)) < RTOL, key
        assert np.max(np.abs(got[key] / gen[key] - 1)) < 1e-11, key
    Nls2 = dict(Nls, gg=Nls['gg'] * np.linspace(1., 1.1, len(g['Llls'])))     # non-constant N_gg: general route only
    a, b = fisher_bias(args[0], args[1], Cls, Nls2, *args[4:]), fisher_bias(args[0], args[1], Cls, Nls2, *args[4:], fused=False)
    assert np.array_equal(a['Fisher'], b['Fisher'])
    ora = ofi.fisher_bias(*args)
    assert np.max(np.abs(got['Fisher'] / ora['Fisher'] - 1)) < RTOL
    assert np.allclose(got['iFisher'], got['iFisher'].T, rtol=1e-12)
    assert np.all(np.linalg.eigvalsh(got['Fisher']) > 0)
    assert np.array_equal(dp_DL('b1', 0.8, 3.0, np.array([1., 2., 3.])), ofi.dp_DL('b1', 0.8, 3.0, np.array([1., 2., 3.])))
    with pytest.raises(ValueError):
        dp_DL('om', 0.8, 3.0, np.ones(3))


def test_fisher_accum_properties(ctx):
    """Random SPD covariances up to 20 x 20 at 4951 multipoles vs numpy; l-shards add up (what the multi-GPU sum relies on)."""
    from lbgcmb_b200.fisher import fisher_accum
    rng = np.random.default_rng(11)
    for ns, npar in ((1, 1), (3, 2), (6, 1), (9, 4), (20, 8)):
        n = 4951 if ns <= 6 else 600
        A = rng.normal(size=(n, ns, ns))
        scale = 10. ** rng.uniform(-8, -2, size=(n, ns))
        cov = (A @ np.transpose(A, (0, 2, 1)) + ns * np.eye(ns)) * scale[:, :, None] * scale[:, None, :]
        dD = rng.normal(size=(n, npar, ns)) * scale[:, None, :]
        dl = rng.normal(size=(n, ns)) * scale
        w = rng.integers(1, 200, n).astype(float)
        F, b, Q = fisher_accum(cov, dD, w, dDelta=dl, per_ell=True)
        icov = np.linalg.inv(cov)
        Qr = np.einsum('lpi,lij,lqj->lpq', dD, icov, dD)
        Fr = np.einsum('l,lpq->pq', w, Qr)
        br = np.einsum('l,lpi,lij,lj->p', w, dD, icov, dl)
        assert np.max(np.abs(Q - Qr) / np.abs(Qr).max(axis=(1, 2), keepdims=True)) < 1e-9
        assert np.max(np.abs(F - Fr)) / np.abs(Fr).max() < RTOL
        assert np.max(np.abs(b - br)) / np.abs(br).max() < 1e-9
        h = n // 2
        F1, _ = fisher_accum(cov[:h], dD[:h], w[:h])
        F2, _ = fisher_accum(cov[h:], dD[h:], w[h:])
        assert np.max(np.abs((F1 + F2) / F - 1)) < 1e-12
    F0, b0 = fisher_accum(np.zeros((0, 3, 3)), np.zeros((0, 2, 3)), np.zeros(0), dDelta=np.zeros((0, 3)))
    assert not F0.any() and not b0.any()
    with pytest.raises(ValueError):
        fisher_accum(np.zeros((2, 40, 40)), np.zeros((2, 1, 40)), np.ones(2))


def test_mqw_fisher_golden(ctx, g):
    """mqw.Fisher on 20 shells x 4941 multipoles vs the reference's percent errors; Fisher matrix vs the oracle."""
    from lbgcmb_b200 import mqw, pmh
    from oracle import fisher as ofi
    from oracle.make_golden import fisher_tracers
    T = fisher_tracers()
    Pk = pmh.get_PkInterps(pmh.PkTable(g['k'], g['plin']))
    zs = np.arange(2.0 + 0.05, 4.0 + 0.05, 0.1)
    out = mqw.Fisher(Pk, g['mqw_L'], 50., 2000., T['mqw_ps'], T['mqw_bs'], T['mqw_pp'], T['mqw_bp'], dz=0.1, fsky=0.35, zmin=2.0,
                     zmax=4.0, fover=0.1, percentiles=list(zs))
    out = np.array([float(np.ravel(x)[0]) for x in out])
    assert out.shape == g['mqw_out'].shape == (40,)
    assert np.max(np.abs(out / g['mqw_out'] - 1)) < 1e-9          # sqrt(diag(inv(F))) of a 20 x 20 matrix
    assert np.max(np.abs(mqw.last['cgg'] / g['mqw_cgg'] - 1)) < RTOL
    F, Fii, S = ofi.mqw_fisher(g['mqw_cgg'], g['mqw_L'].astype(float), mqw.last['bins'], 0.35)
    assert np.max(np.abs(mqw.last['Fisher'] / F - 1)) < RTOL
    assert np.max(np.abs(mqw.last['Fii'] / Fii - 1)) < RTOL
    assert np.max(np.abs(mqw.last['S'] / S - 1)) < RTOL


def test_rho_L_golden(ctx, g):
    from lbgcmb_b200 import pmh, yu
    from oracle.make_golden import fisher_tracers
    T = fisher_tracers()
    Pk = pmh.get_PkInterps(pmh.PkTable(g['k'], g['plin']))
    rho = yu.rho_L(Pk, g['yu_L'], g['yu_ckk'], T['yu'])
    assert np.max(np.abs(rho / g['yu_rho'] - 1)) < RTOL
    assert np.all((rho > 0) & (rho < 1))
