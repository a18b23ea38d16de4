"""GPU parity: Eisenstein-Hu transfer function (csrc/eh.cu) vs the golden vectors written by the COMPILED reference
(EisensteinHu/EisensteinAndHu.c -> oracle/_ref/libeh_ref.so, float32: tolerance 2e-6) and vs the float64 numpy oracle
(oracle/eh.py: tolerance 1e-12); linear-table builder vs the oracle (1e-12) and as a working P(k) provider."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_transfer_vs_reference_vectors_and_oracle(ctx, golden_dir):
    from lbgcmb_b200.eisenstein_hu import TFmdm
    from oracle import eh
    g = np.load(os.path.join(golden_dir, 'eh.npz'))
    k = g['k']
    for i, case in enumerate(g['cases']):
        args = tuple(case[:3]) + (int(case[3]),) + tuple(case[4:])
        tf = TFmdm(*args)
        cb = tf.onek_hmpc(k)
        assert np.max(np.abs(cb / g['tf_cb_%d' % i] - 1)) < 2e-6
        assert np.max(np.abs(tf.tf_cbnu / g['tf_cbnu_%d' % i] - 1)) < 2e-6
        c = eh.set_cosm(*args)
        ocb, ocbnu = eh.onek_hmpc(c, k)
        assert np.max(np.abs(cb / ocb - 1)) < 1e-12 and np.max(np.abs(tf.tf_cbnu / ocbnu - 1)) < 1e-12
        assert abs(tf.growth_to_z0 / c['growth_to_z0'] - 1) < 1e-14
        assert np.max(np.abs(tf.onek_mpc(k * args[5]) / cb - 1)) < 1e-14
    assert isinstance(TFmdm(*args).onek_hmpc(0.1), float)
    with pytest.raises(ValueError):
        TFmdm(0.3, 0.05, 0.0, 1, 0.7, -0.7, 0.0)
    with pytest.raises(ValueError):
        TFmdm(0.3, 0.05, 0.0, 1, 0.7, 0.7, -1.5)


def test_linear_table_vs_oracle_and_as_provider(ctx):
    from lbgcmb_b200 import eisenstein_hu as ehm, pmh, lensing
    from lbgcmb_b200.synthetic import PARAMS
    from oracle import eh
    k = np.exp(np.linspace(np.log(1.5e-5), np.log(100.), 2000))
    p = ehm.linear_power(k, PARAMS['om_m'], PARAMS['om_b'], 1. - PARAMS['om_m'], PARAMS['h_100'], PARAMS['ns'], PARAMS['sig_8'])
    ref = eh.linear_power(k, (PARAMS['om_m'], PARAMS['om_b'], 0.0, 1, 1. - PARAMS['om_m'], PARAMS['h_100'], 0.0), PARAMS['ns'], PARAMS['sig_8'])
    assert np.max(np.abs(p / ref - 1)) < 1e-12
    # sigma8 of the table really is sigma8 (independent quadrature), and it drives the Limber path
    x = 8. * k
    w = 3. * (np.sin(x) - x * np.cos(x)) / x ** 3
    s8 = np.sqrt(np.trapezoid(p * k ** 3 / (2 * np.pi ** 2) * w * w, np.log(k)))
    assert abs(s8 / PARAMS['sig_8'] - 1) < 1e-6
    Pk = pmh.get_PkInterps(pmh.PkTable(k, p))
    ckk = lensing.Ckk(Pk, np.array([100., 1000.]))
    assert np.all(ckk > 0) and ckk[0] > ckk[1]
