"""GPU parity of the batched experiment x delensing x dilution sweep (csrc/fisher.cu snr_grid_kernel, lbg_snr_grid; snr.snr_grid,
snr.dilution_sweep) against the per-design route of the reference's driver, snr.py:107-133: var_Ckg (Ckg.py:61-87, incl. the
`rho=` branch) followed by snr (snr.py:21-25), both through the mirrors that tests/test_variances_gpu.py and tests/test_fisher_gpu.py pin
to golden vectors of the UNMODIFIED reference.  Given the same N_kk curves the grid call is bit-identical; the driver-level sweep
(which may take two experiments of one beam through the multi-level N0 kernel) agrees to 1e-10."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10
DF = np.array([1.e-5, 0.001, 0.005, 0.01, 0.05, 0.1, 0.5, 1.0, 5.0, 1.e1, 5.e1, 3.e3])        # snr.py:108


@pytest.fixture(scope='module')
def setup(ctx, golden_dir):
    from lbgcmb_b200 import pmh
    from oracle.make_golden import variance_inputs          # analytic input callables only (no oracle arithmetic)
    g = np.load(os.path.join(golden_dir, 'variances.npz'))
    T = variance_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(g['k'], g['plin']))
    return Pk, T, g


def test_grid_equals_var_Ckg_then_snr_bit_for_bit(setup):
    from lbgcmb_b200 import snr as S
    from lbgcmb_b200.Cgg import Cgg, Ngg
    from lbgcmb_b200.Ckg import Ckg, var_Ckg
    from lbgcmb_b200.lensing import Ckk
    Pk, T, g = setup
    L = g['Llls'].astype(np.float64)
    nm = np.linspace(1., 3., L.size)
    nkk = np.array([g['nkk'], 2.5 * g['nkk'], 0.3 * g['nkk']])
    fsky = np.array([0.4, 1.0, 0.2])
    rhos = [None, T['rho']]
    z = (T['zmin'], T['zmax'], T['pz'], T['bz'])
    ckg, ckk, cgg = Ckg(Pk, L, *z, zeff=True), Ckk(Pk, L), Cgg(Pk, L, *z, zeff=True)
    ngg = np.array([Ngg(L, T['zmin'], T['zmax'], T['pz'], f * T['nbar'])[0] for f in DF])
    rho = np.array([np.zeros_like(L), T['rho'](L)])
    lmax = float(L[L.size * 2 // 3])
    got = S.snr_grid(L, ckg, ckk, cgg, nkk, fsky, rho, ngg, nm, lmax=lmax)
    assert got.shape == (3, 2, DF.size) and np.all(got > 0)
    for e in range(3):
        for r in range(2):
            for d in (0, 3, 7, 11):
                v = var_Ckg(Pk, None, None, L, *z, DF[d] * T['nbar'], fsky[e], nkk=nkk[e], rho=rhos[r])
                assert got[e, r, d] == S.snr(L, ckg, v, nm, lmax=lmax)
    assert np.all(got[:, 1] >= got[:, 0])                       # delensing lowers the variance, never the S/N
    assert np.all(np.diff(got, axis=2) > 0)                     # denser samples, more S/N
    one = S.snr_grid(L, ckg, ckk, cgg, nkk[1], fsky[1:2], rho[1], ngg[4:5], nm, lmax=lmax)          # 1-d inputs for a single design
    assert one.shape == (1, 1, 1) and one[0, 0, 0] == got[1, 1, 4]
    assert np.array_equal(S.snr_grid(L, ckg, ckk, cgg, nkk, fsky, rho, ngg, nm, lmax=1.0), np.zeros((3, 2, DF.size)))    # nothing below lmax
    with pytest.raises(ValueError):
        S.snr_grid(L, ckg, ckk, cgg, nkk, fsky[:2], rho, ngg, nm)
    with pytest.raises(ValueError):
        S.snr_grid(L, ckg, ckk, cgg, nkk, -fsky, rho, ngg, nm)


def test_dilution_sweep_equals_the_reference_loop_nest(setup, cl_tables):
    """snr.dilution_sweep for the reference's bolometers (SO and CMBS4 share a beam -> one multi-level N0 pass) against the loop nest of
    snr.py:107-133 run through the per-function mirrors."""
    from lbgcmb_b200 import snr as S
    from lbgcmb_b200.Ckg import Ckg, var_Ckg
    from lbgcmb_b200.Nkk import Nkk
    from lbgcmb_b200.bolometers import bolometers
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    Pk, T, g = setup
    li, ni = cl_interps_from_tables(*cl_tables)
    L = np.unique(np.round(np.logspace(np.log10(50.), np.log10(4000.), 24))).astype(np.float64)
    nm = np.ones_like(L)
    z = (T['zmin'], T['zmax'], T['pz'], T['bz'])
    exps = {k: bolometers[k] for k in ('Planck', 'SO', 'CMBS4')}
    dfs = DF[[0, 5, 7, 11]]
    got, names = S.dilution_sweep(Pk, li, ni, L, nm, *z, T['nbar'], exps, [None, T['rho']], dfs, Lmax=2000.)
    assert names == ['Planck', 'SO', 'CMBS4'] and got.shape == (3, 2, 4)
    ckg = Ckg(Pk, L, *z, zeff=True)
    for e, name in enumerate(names):
        b = exps[name]
        nkk = Nkk(li, ni, L, ['TT', 'TE', 'EE', 'EB'], b['thetab'], b['DeltaT'], iterative=b['iterative'], pickleit=False)
        for r, rho in enumerate([None, T['rho']]):
            for d, f in enumerate(dfs):
                v = var_Ckg(Pk, li, ni, L, *z, f * T['nbar'], b['fsky'], nkk=nkk, iterative=b['iterative'], rho=rho)
                want = S.snr(L, ckg, v, nm, lmax=2000.) / np.sqrt(b['fsky'])
                assert abs(got[e, r, d] / want - 1) < RTOL
    assert got[2].min() > got[0].max() * 0 and np.all(got[2] > got[0])           # CMB-S4 beats Planck at every dilution
