"""CLASS C_l file readers: lbgcmb_b200/prep_class.py against the UNMODIFIED reference prep_class.py
(build container) and, on the GPU, against scipy's FITPACK spline and the table route of Nkk."""
import os
import re
import sys
import types

import numpy as np
import pytest


def write_class_files(root):
    """dat/RunPB/RunPB02_cl{,_lensed}.dat under `root` from the synthetic tables, in CLASS layout (l = 2.., l(l+1)/2pi C_l)."""
    from lbgcmb_b200.synthetic import cmb_cl_tables
    ell, lens, nolens = cmb_cl_tables()
    d = os.path.join(root, 'dat', 'RunPB')
    os.makedirs(d, exist_ok=True)
    paths = []
    for name, tab, lmax in (('RunPB02_cl_lensed.dat', lens, 6000), ('RunPB02_cl.dat', nolens, 6500)):
        l = ell[2:lmax + 1]
        pre = l * (l + 1.) / (2. * np.pi)
        tt, ee, bb, te = (tab[i][2:lmax + 1] * pre for i in range(4))
        pp = 1e-7 / l ** 4. * pre
        np.savetxt(os.path.join(d, name), np.c_[l, tt, ee, te, bb, pp, 0. * l, 0. * l], fmt='%.17e')
        paths.append(os.path.join(d, name))
    return paths, (ell, lens, nolens)


@pytest.mark.reference
def test_tables_match_reference_prep_class(tmp_path, monkeypatch):
    from oracle import refload
    if not refload.available():
        pytest.skip('/root/reference not present (GPU box)')
    refload.setup()
    import pandas as pd
    from lbgcmb_b200 import prep_class
    (lensed, unlensed), _ = write_class_files(str(tmp_path))
    real_read_csv = pd.read_csv

    def read_csv(*a, **kw):                       # pandas >= 3 dropped delim_whitespace (prep_class.py:21,39)
        if kw.pop('delim_whitespace', False):
            kw['sep'] = r'\s+'
        return real_read_csv(*a, **kw)
    monkeypatch.setattr(pd, 'read_csv', read_csv)
    path = os.path.join(refload.REF, 'prep_class.py')
    src = re.sub(r'^(\s*)print\s+(".*")\s*$', r'\1print(\2)', open(path).read(), flags=re.M)   # python-2 prints in __main__
    mod = types.ModuleType('prep_class_ref')
    mod.__file__ = path
    exec(compile(src, path, 'exec'), mod.__dict__)
    monkeypatch.chdir(tmp_path)                   # the reference reads dat/RunPB/... relative to the cwd
    rl, rn = mod.prep_classCls()
    gl, gn = prep_class.prep_classCls()           # same default paths
    for x in ('TT', 'EE', 'TE', 'BB'):
        for got, ref in ((gl[x], rl[x]), (gn[x], rn[x])):
            rx, ry = ref._data[0], ref._data[1]
            assert np.array_equal(got.x, rx)
            assert np.allclose(got.y, ry, rtol=1e-15, atol=0)
    d = prep_class.read_class(lensed)
    assert np.allclose(d['kk'], 0.25 * (d['ell'] * (d['ell'] + 1.)) ** 2. * d['PP'], rtol=0, atol=0)


def test_class_cl_tables_layout(tmp_path):
    from lbgcmb_b200 import prep_class
    (lensed, unlensed), (ell, lens, nolens) = write_class_files(str(tmp_path))
    e, lt, ut = prep_class.class_cl_tables(lensed, unlensed)
    assert e[0] == 0 and e[-1] == 6000 and lt.shape == ut.shape == (4, 6001)
    assert np.all(lt[:, :2] == 0) and np.all(ut[:, :2] == 0)
    assert np.allclose(lt[:, 2:], lens[:, 2:6001], rtol=1e-15, atol=0)
    assert np.allclose(ut[:, 2:], nolens[:, 2:6001], rtol=1e-15, atol=0)
    bad = tmp_path / 'short.dat'
    np.savetxt(bad, np.ones((5, 3)))
    with pytest.raises(ValueError):
        prep_class.read_class(str(bad))


@pytest.mark.gpu
def test_class_splines_and_n0_on_gpu(tmp_path):
    from scipy.interpolate import InterpolatedUnivariateSpline
    from lbgcmb_b200 import prep_class, Nkk
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    (lensed, unlensed), (ell, lens, nolens) = write_class_files(str(tmp_path))
    gl, gn = prep_class.prep_classCls(lensed, unlensed)
    x = np.concatenate([np.linspace(2., 6000., 4001), [1.2, 6003.7]])      # two extrapolated points (ext = 0)
    for f in (gl['TT'], gl['BB'], gn['TE'], gn['EE']):
        ref = InterpolatedUnivariateSpline(f.x, f.y, ext=0)(x)
        assert np.max(np.abs(f(x) - ref) / np.max(np.abs(ref))) < 1e-12
    Ls = np.array([50, 1003, 5000])
    li, ni = cl_interps_from_tables(*prep_class.class_cl_tables(lensed, unlensed))
    li0, ni0 = cl_interps_from_tables(ell, lens, nolens)
    a = Nkk.Nkk(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 7.0, iterative=False, pickleit=False)
    b = Nkk.Nkk(li0, ni0, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 7.0, iterative=False, pickleit=False)
    assert np.max(np.abs(a / b - 1)) < 1e-9          # tables cut at l = 6000 instead of the synthetic lmax; l(l+1)/2pi round trip
