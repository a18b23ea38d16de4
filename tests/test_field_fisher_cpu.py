"""The field-form Fisher restatement (oracle/fisher.py field_fisher) pinned to the UNMODIFIED reference: with fields (kappa, g),
spectra (kk, kg, gg) and w = nmodes (2L+1) fsky it reproduces the interloper.py:200-253 Fisher matrix of tests/golden/fisher.npz."""
import os

import numpy as np


def test_oracle_field_form_equals_reference_fisher(golden_dir):
    from oracle.fisher import field_fisher
    g = np.load(os.path.join(golden_dir, 'fisher.npz'))
    L, nm, fsky = g['Llls'], g['nmodes'], float(g['fsky'])
    s8, b1 = float(g['fid_sig8']), float(g['fid_b1'])
    tot = np.array([g['C_kk'] + g['N_kk'], g['C_kg'], g['C_gg'] + g['N_gg']])
    z = np.zeros_like(g['C_kk'])
    dC = np.array([[2. / s8 * g['C_kk'], 2. / s8 * g['C_kg'], 2. / s8 * g['C_gg']], [z, g['C_kg'] / b1, 2. * g['C_gg'] / b1]])
    w = nm * (2. * L + 1.) * fsky * (L < float(g['LllMax']))
    F, _ = field_fisher(2, [(0, 0), (0, 1), (1, 1)], tot, dC, w)
    assert np.max(np.abs(F / g['Fisher'] - 1)) < 1e-10
