"""GPU parity of the variance wrappers and small helpers of SURVEY.md 8(a) rows a3, a5, a6, a7, a9 -- Cgg.var_Cgg (Cgg.py:70-83),
Ckg.var_Ckg incl. `rho=` / `modesample_lim` / `ngg=` / `nkk=None` (Ckg.py:60-87), lensing.var_Ckk (cmb/lensing.py:54-66),
prep_camb.Clxy (:128-137), pmh.Pmm / Pmh / Phh (:26-47) -- against golden vectors written by the UNMODIFIED reference
(tests/golden/variances.npz; oracle/make_golden.py make_variances).  Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


@pytest.fixture(scope='module')
def g(golden_dir):
    return np.load(os.path.join(golden_dir, 'variances.npz'))


@pytest.fixture(scope='module')
def setup(ctx, g):
    from lbgcmb_b200 import pmh
    from oracle.make_golden import variance_inputs          # analytic input callables only (no oracle arithmetic)
    T = variance_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(g['k'], g['plin']))
    return Pk, T


def _rel(a, b):
    return float(np.max(np.abs(np.asarray(a) / np.asarray(b) - 1)))


def test_var_Cgg(g, setup):
    from lbgcmb_b200.Cgg import var_Cgg, Cgg
    Pk, T = setup
    L = g['Llls'].astype(np.float64)
    a = (L, T['zmin'], T['zmax'], T['pz'], T['bz'], T['nbar'], T['fsky'])
    assert _rel(var_Cgg(*a, Pk_interps=Pk), g['var_cgg']) < RTOL                       # cgg=None: zeff=True spectrum inside
    assert _rel(var_Cgg(*a, Pk_interps=Pk, samplevar_lim=True), g['var_cgg_sv']) < RTOL
    cgg = Cgg(Pk, L, T['zmin'], T['zmax'], T['pz'], T['bz'], zeff=False)
    assert _rel(cgg, g['cgg_full']) < RTOL
    assert _rel(var_Cgg(*a, cgg=cgg), g['var_cgg_given']) < RTOL
    assert np.array_equal(var_Cgg(*a, cgg=g['cgg_full']), g['var_cgg_given'])          # given spectrum: same three lines, same bits


def test_var_Ckg_branches(g, setup):
    from lbgcmb_b200.Ckg import var_Ckg
    Pk, T = setup
    L, nkk = g['Llls'].astype(np.float64), g['nkk']
    vk = lambda **kw: var_Ckg(Pk, None, None, L, T['zmin'], T['zmax'], T['pz'], T['bz'], T['nbar'], T['fsky'], nkk=nkk, **kw)
    assert _rel(vk(), g['var_ckg']) < RTOL
    assert _rel(vk(zeff=False), g['var_ckg_full']) < RTOL
    assert _rel(vk(rho=T['rho']), g['var_ckg_rho']) < RTOL                             # delensed C_kk (Ckg.py:69-71)
    assert _rel(vk(modesample_lim=True), g['var_ckg_msl']) < RTOL
    assert _rel(vk(ngg=3.0e-9 * np.ones_like(L)), g['var_ckg_ngg']) < RTOL
    assert np.all(vk(rho=T['rho']) < vk())                                             # delensing can only lower the variance


def test_var_Ckg_runs_Nkk_when_nkk_is_None(g, setup):
    """nkk=None: the wrapper computes N0 itself with the thetab / DeltaT / iterative it was given (Ckg.py:80-82)."""
    from lbgcmb_b200.Ckg import var_Ckg
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    from lbgcmb_b200.synthetic import cmb_cl_tables
    Pk, T = setup
    li, ni = cl_interps_from_tables(*cmb_cl_tables(5300))
    got = var_Ckg(Pk, li, ni, g['L2'], T['zmin'], T['zmax'], T['pz'], T['bz'], T['nbar'], T['fsky'], thetab=1.4, DeltaT=7.0,
                  iterative=True, pickleit=False)
    assert _rel(got, g['var_ckg_nkknone']) < RTOL


def test_var_Ckk(g, setup):
    from lbgcmb_b200.lensing import var_Ckk
    Pk, T = setup
    L, nkk = g['Llls'].astype(np.float64), g['nkk']
    assert _rel(var_Ckk(L, T['fsky'], nkk, Pk_interps=Pk), g['var_ckk']) < RTOL
    assert _rel(var_Ckk(L, T['fsky'], nkk, Pk_interps=Pk, samplevar_lim=True), g['var_ckk_sv']) < RTOL


def test_Clxy(ctx, g):
    """C_l + detector noise on integer and fractional multipoles; TE carries no noise; unknown spectrum -> KeyError / ValueError."""
    from lbgcmb_b200.prep_camb import Clxy, cl_interps_from_tables
    from lbgcmb_b200.synthetic import cmb_cl_tables
    li, _ = cl_interps_from_tables(*cmb_cl_tables(5300))
    x = g['clxy_x']
    for a in ('TT', 'EE', 'BB', 'TE'):
        got = Clxy(li, x.copy(), a, 1.4, 7.0)
        assert np.max(np.abs(got - g['clxy_' + a])) <= RTOL * np.max(np.abs(g['clxy_' + a])), a
    assert np.array_equal(Clxy(li, x.copy(), 'TE', 1.4, 7.0), li['TE'](x))
    with pytest.raises(KeyError):
        Clxy(li, x, 'TB', 1.4, 7.0)


def test_Pmm_Pmh_Phh(g, setup):
    from lbgcmb_b200.pmh import Pmm, Pmh, Phh
    Pk, T = setup
    k, z = g['pk_k'], float(g['pk_z'])
    assert _rel(Pmm(Pk, k, z, type='linear'), g['pmm_lin']) < RTOL
    assert _rel(Pmm(Pk, k, z), g['pmm_nlin']) < RTOL
    assert _rel(Pmh(Pk, T['bz'], k, z), g['pmh']) < RTOL
    assert _rel(Phh(Pk, T['bz'], k, z), g['phh']) < RTOL
    assert _rel(Phh(Pk, T['bz'], k, z, bz2=T['bz2']), g['phh2']) < RTOL
    with pytest.raises(ValueError):
        Pmm(Pk, k, z, type='halo')
