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
