"""Build container only (needs /root/reference and its data files): the input-provider mirrors return what the UNMODIFIED
reference functions return -- dropouts/schechter/get_pz.py, get_shot.py, gen_pz.peakz, dropouts/bz/py/get_bz.py,
schmittfull_nz.py / schmittfull_bz.py, whitebook_pz.py, growth_rate.py."""
import os
import sys

import numpy as np
import numpy.ma as ma
import pytest

pytestmark = pytest.mark.reference


@pytest.fixture(scope='module')
def ref():
    from oracle import refload
    if not refload.available():
        pytest.skip('/root/reference not present (GPU box)')
    refload.setup()
    os.environ['LBGCMB'] = refload.REF
    for sub in ('dropouts', 'dropouts/bz/py', 'dropouts/schechter'):
        p = os.path.join(refload.REF, sub)
        if p not in sys.path:
            sys.path.insert(0, p)
    return refload


def test_dropout_providers(ref):
    from lbgcmb_b200.schechter.get_pz import get_pz
    from lbgcmb_b200.schechter.get_shot import get_shot
    from lbgcmb_b200.schechter.gen_pz import peakz
    from lbgcmb_b200.get_bz import bz_callmodel, drop_bz
    rpz, rshot = ref.load('schechter.get_pz'), ref.load('schechter.get_shot')
    rbz = ref.load('get_bz')
    zs = np.arange(0., 8., 0.013)
    for band, mlim in (('BX', 25.5), ('u', 24.6), ('g', 25.8), ('r', 25.8)):
        assert np.array_equal(get_pz(band)(zs), rpz.get_pz(band)(zs))
        assert get_shot(band, mlim) == rshot.get_shot(band, mlim)
        assert np.array_equal(bz_callmodel(zs, mlim), rbz.bz_callmodel(zs, mlim))
    ms, ns = get_shot('qso', None)
    rm, rn = rshot.get_shot('qso', None)
    assert np.array_equal(ms, rm) and np.array_equal(ns, rn)
    assert peakz(get_pz('u')) == 2.99           # gen_pz.py itself imports modules that are not in the tree; :37-43 restated
    a, b = drop_bz(), rbz.drop_bz()
    assert np.array_equal(a.zs, b.zs) and np.array_equal(a.ms, b.ms) and np.array_equal(a.bs.filled(-1), b.bs.filled(-1))
    assert np.array_equal(a.el.filled(-1), b.el.filled(-1)) and np.array_equal(a.eu.filled(-1), b.eu.filled(-1))
    for z, m in ((a.zs[0], a.ms[0]), (a.zs[1] + 0.01, a.ms[-1] - 0.01), (0.5 * (a.zlo + a.zhi), 0.5 * (a.mlo + a.mhi))):
        for x, y in zip(a.get_bzm(z, m), b.get_bzm(z, m)):
            assert x.shape == y.shape and np.array_equal(ma.filled(x, -1), ma.filled(y, -1))
    with pytest.raises(UserWarning):
        a.get_bzm(9.0, 25.0)


def test_ss17_and_whitebook_providers(ref):
    from lbgcmb_b200 import schmittfull_nz as nz, schmittfull_bz as bz, whitebook_pz as wb, growth_rate as gr
    cwd = os.getcwd()
    os.chdir(ref.REF)                                  # the reference reads dat/... relative to the cwd
    try:
        rnz, rbz, rwb, rgr = ref.load('schmittfull_nz'), ref.load('schmittfull_bz'), ref.load('whitebook_pz'), ref.load('growth_rate')
        ns, ps, bs, ss = nz.get_ss17_samples(nolsst=True)
        rns, rps, rbs, rss = rnz.get_ss17_samples(nolsst=True)
    finally:
        os.chdir(cwd)
    zs = np.array([0.05, 0.3, 0.7, 1.0, 1.5, 2.5])
    assert ss == rss and np.array_equal(ns, rns)
    for p, rp, b, rb in zip(ps, rps, bs, rbs):
        assert np.array_equal(p(zs), rp(zs))           # includes the late-binding closure: every p(z) is the last row's top-hat
        assert np.allclose(b(zs), rb(zs), rtol=1e-14)
    assert np.array_equal(ps[0](zs), wb.const_pz(zs, 0.6, 1.9))
    aa = np.array([0.2, 0.5, 1.0])
    assert np.allclose(gr.growth_factor(aa), rgr.growth_factor(aa), rtol=1e-14)
    for f in ('whitebook_pz', 'lsst_chang_pz'):
        assert np.allclose(getattr(wb, f)(zs, 25.3), getattr(rwb, f)(zs, 25.3), rtol=1e-14)
    assert wb.Ng(25.3) == rwb.Ng(25.3) and np.array_equal(wb.whitebook_dpdz(zs, 25.3), rwb.whitebook_dpdz(zs, 25.3))
    with pytest.raises(ValueError):
        bz.desi_bgs(np.array([0.7]))
