"""CPU: host-side input providers (whitebook_pz.py mirror) against closed forms and their own normalisation."""
import numpy as np
import pytest


def test_whitebook_pz_providers():
    from lbgcmb_b200 import whitebook_pz as wb
    zs = np.arange(0., 30., 1e-3)
    assert abs(np.trapezoid(wb.whitebook_pz(zs, 25.3), zs) - 1.) < 1e-6            # normalised (science book 3.8)
    h = 1e-6
    assert abs((wb.whitebook_pz(0.7 + h, 25.3) - wb.whitebook_pz(0.7 - h, 25.3)) / (2 * h) / wb.whitebook_dpdz(0.7, 25.3) - 1) < 1e-6
    assert wb.const_pz(0.5) == 1.0 and wb.const_pz(1.5) == 0.0
    assert np.array_equal(wb.const_pz(np.array([0.5, 2.0]), 0.0, 1.0), [1.0, 0.0])
    assert abs(wb.Ng(25.3) / (42.9 * 0.88 * 10 ** (0.359 * 0.3) * 3600.) - 1) < 1e-14
    z = np.arange(0., 10., 0.01)
    for src in (False, True):
        assert abs(np.sum(wb.lsst_chang_pz(z, 25.3, src)) * 0.01 - 1.) < 1e-12
    with pytest.raises(ValueError):
        wb.lsst_chang_pz(0.5, ilim=22.0)


def test_result_writers(tmp_path):
    """rho/*.txt (yu.py:113) and dat/result4interloper_*.json (interloper.py:271-275) layouts."""
    import json
    from lbgcmb_b200.interloper import save_result
    from lbgcmb_b200.yu import save_rho
    L, rho = np.array([50, 100, 200]), np.array([0.5, 0.6, 0.7])
    save_rho(str(tmp_path / 'a_b.txt'), L, rho)
    back = np.loadtxt(str(tmp_path / 'a_b.txt'))
    assert np.allclose(back[:, 0], L) and np.allclose(back[:, 1], 1. - rho ** 2.)
    res = {'fid_sig8': np.float64(0.3), 'fid_b1': 3.96, 'biased_sig8': 0.31, 'biased_b1': 3.9, 'iFisher': np.eye(2) * 1e-4}
    save_result(str(tmp_path / 'r.json'), 2.99, 'u', 0.06, res)
    d = json.load(open(str(tmp_path / 'r.json')))
    assert sorted(d) == ['biased_b1', 'biased_sig8', 'fid_b1', 'fid_sig8', 'iFisher', 'intfrac', 'peakz'] and d['iFisher'][1][1] == 1e-4
