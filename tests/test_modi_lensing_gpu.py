"""GPU parity: the modi-flavour Limber path (mirrors of modi/cosmology.py and modi/lensing.py over csrc/limber.cu) vs golden
vectors written by the reference's own modi/cosmology.py + modi/lensing.py (tests/golden/modi_lensing.npz,
oracle/refload.load_modi) and the numpy oracle.  Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10

dndz = lambda z: z ** 2. * np.exp(-(z / 0.5) ** 1.0)
dndz2 = lambda z: z ** 1.5 * np.exp(-(z / 0.4) ** 1.2)
bz = lambda z: 1.438 + 0. * z


@pytest.fixture(scope='module')
def g(golden_dir):
    return np.load(os.path.join(golden_dir, 'modi_lensing.npz'))


@pytest.fixture(scope='module')
def cosmo(g):
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200.synthetic import PARAMS
    return Cosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=g['k'], plin=g['plin'])


def test_background_and_pk_rows(ctx, g, cosmo):
    assert np.array_equal([cosmo.Dgrow(z=z) for z in g['zchk']], g['Dchk'])         # same Romberg, same spline
    assert np.allclose([cosmo.chia(z=z) for z in g['zchk']], g['chichk'], rtol=1e-14)
    k, p = cosmo.pkanlin(z=1.0)
    assert np.max(np.abs(p / g['pknl_z1'] - 1)) < RTOL
    assert np.max(np.abs(cosmo.pkanlin(z=5.0)[1] / g['pklin_z5'] - 1)) < RTOL      # z > 4 -> linear (modi/cosmology.py:171)
    assert np.max(np.abs(cosmo.pkalin(z=5.0)[1] / g['pklin_z5'] - 1)) < RTOL


def test_lensingx_golden(ctx, g, cosmo):
    from lbgcmb_b200.modi.lensing import LensingX, PkZ, prefactor
    lx = LensingX(z=1.0, dndz=dndz, cosmo=cosmo, l=g['l'])
    for name in ('kerg', 'kercmb', 'dggwtsdN', 'dkgwtsdN', 'kkwts', 'chis'):
        assert np.allclose(getattr(lx, name), g[name], rtol=1e-13), name
    assert abs(lx.norm / g['norm'] - 1) < 1e-13
    mmz, hmz, hhz = PkZ(cosmo), PkZ(cosmo, bias=bz, power=1), PkZ(cosmo, bias=bz, power=2)
    got = {'clkkz': lx.cl_kkz(mmz), 'clggz': lx.clz(hhz, auto=True), 'clkgz': lx.clz(hmz, auto=False),
           'clggzeff': lx.cl_zeff(hhz(1.0), auto=True), 'clkgzeff': lx.cl_zeff(hmz(1.0), auto=False),
           'clbitracer': lx.cl_zeff_bitracer(hhz(1.0), dndz2)}
    for name, v in got.items():
        rel = np.max(np.abs(v / g[name] - 1))
        assert rel < RTOL, '%s rel=%.3e' % (name, rel)
    # a foreign callable (the reference's own lambdas) goes through the host-sampled rows and gives the same numbers
    lam = lambda z: [cosmo.klin, cosmo.pkanlin(z=z)[1] * bz(z)]
    assert np.max(np.abs(lx.clz(lam, auto=False) / got['clkgz'] - 1)) < 1e-13
    assert lx.kmesh.shape == (g['l'].size, 100) and lx.kk_kmesh.shape == (g['l'].size, 2000)
    assert prefactor(2., 0) == 1.0 and prefactor(np.array([2.]), 2)[0] == 6. / (2. * np.pi)
    with pytest.raises(ValueError):
        LensingX(z=1.0, dndz=dndz)


def test_lensingx_ggnorm_and_default_l(ctx, g, cosmo, capsys):
    from lbgcmb_b200.modi.lensing import LensingX
    lx0 = LensingX(z=1.0, dndz=dndz, cosmo=cosmo, nz=20)
    lx = LensingX(z=1.0, dndz=dndz, cosmo=cosmo, ggnorm=2.0 * lx0.norm, nz=20)
    assert 'Overriding estimated norm' in capsys.readouterr().out
    assert lx.l.size == 500 and lx.kknz == 400
    assert np.allclose(lx.kerg * 2.0, lx0.kerg, rtol=1e-14)
