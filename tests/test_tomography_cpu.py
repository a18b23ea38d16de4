"""Host logic of lbgcmb_b200/tomography.py that needs no GPU: the vectorised comoving distance against the `quad` of the Cosmology
mirror (modi/cosmology.py:134-150, itself pinned to the reference in tests/test_modi_lensing_gpu.py)."""
import numpy as np


def test_background_distance_equals_quad():
    from lbgcmb_b200.modi.cosmology import Cosmology
    from lbgcmb_b200 import tomography as tm
    y = np.ones(1000)                                   # growth table is irrelevant here; skips the 1000 host Romberg integrals
    for M, L in ((0.292, None), (0.31, 0.69), (0.25, 0.7)):
        ref = Cosmology(M=M, L=L, Dplus_table=y)
        fast = tm.Background(M=M, L=L, Dplus_table=y)
        z = np.concatenate([[0.0, 1e-3, 0.05], np.linspace(0.1, 10., 23), [50., 1100.]])
        a, b = ref.chia(z=z), fast.chia(z=z)
        assert b.shape == z.shape and b[0] == 0.0
        assert np.max(np.abs(b[1:] / a[1:] - 1)) < 1e-13
        assert abs(fast.chia(z=1100.) / ref.chia(z=1100.) - 1) < 1e-13 and isinstance(fast.chia(z=2.0), float)
        assert abs(fast.chia(a=0.5) / ref.chia(a=0.5) - 1) < 1e-13
    assert abs(tm.omega_hdm(0.06, 0.69) - 0.06 / 93.14 / 0.69 ** 2) < 1e-18
