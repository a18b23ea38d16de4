"""GPU parity: field-level Fisher sum (csrc/fisher.cu field_fisher_kernel, lbg_field_fisher) vs
 (1) the reference-pinned spectra form (interloper.py:42-63 covariance + :223-253 sums, tests/golden/fisher.npz) for fields (kappa, g),
 (2) the dense numpy restatement (oracle/fisher.py field_fisher) on random positive-definite cases up to the 121 fields of
     BASELINE configs[3] (kappa + 6 tracers x 20 bins), full and arrowhead (disjoint-bin) patterns.
Tolerance: relative 1e-10 (FP64)."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def _random_case(rng, n_f, n_ell, n_par, arrow):
    """SPD field covariances per multipole with the given sparsity pattern and random symmetric derivatives on the same pattern."""
    if arrow == 2:                                      # two dense fields (CMB and galaxy lensing, say) x every bin
        spectra = [(0, 0), (0, 1), (1, 1)] + [(h, i) for h in (0, 1) for i in range(2, n_f)] + [(i, i) for i in range(2, n_f)]
    elif arrow:                                         # kappa x every bin, bins mutually uncorrelated
        spectra = [(0, 0)] + [(0, i) for i in range(1, n_f)] + [(i, i) for i in range(1, n_f)]
    else:
        spectra = [(a, b) for a in range(n_f) for b in range(a, n_f)]
    n_spec = len(spectra)
    tot = np.zeros((n_spec, n_ell))
    for l in range(n_ell):
        d = 10. ** rng.uniform(-8, -4, n_f)             # auto spectra over four decades, as C_kk vs C_gg + shot noise
        for s, (a, b) in enumerate(spectra):
            if a == b:
                tot[s, l] = d[a]
            else:
                r = rng.uniform(-1, 1) * (0.4 / np.sqrt(n_f) if arrow else 0.5 / n_f)   # keeps C diagonally dominant after scaling
                tot[s, l] = r * np.sqrt(d[a] * d[b])
    dC = tot[None] * rng.uniform(-2, 2, (n_par, n_spec, 1)) * rng.uniform(0.5, 1.5, (n_par, n_spec, n_ell))
    w = rng.uniform(1, 100, n_ell)
    return spectra, tot, dC, w


@pytest.mark.parametrize('n_f,n_ell,n_par,arrow', [(1, 5, 1, False), (2, 40, 2, False), (7, 33, 3, False), (21, 20, 8, True),
                                                   (32, 9, 4, False), (121, 6, 4, True), (160, 2, 2, True),
                                                   (160, 3, 8, True), (160, 2, 8, 2), (40, 11, 5, 2), (33, 300, 2, True)])
def test_field_fisher_vs_dense_numpy(ctx, n_f, n_ell, n_par, arrow):
    from lbgcmb_b200.fisher import field_fisher
    from oracle.fisher import field_fisher as ref
    rng = np.random.default_rng(1000 * n_f + n_par)
    spectra, tot, dC, w = _random_case(rng, n_f, n_ell, n_par, arrow)
    F, Q = field_fisher(n_f, spectra, tot, dC, w, per_ell=True)
    F0, Q0 = ref(n_f, spectra, tot, dC, w)
    scale = np.sqrt(np.abs(np.einsum('lpp,lqq->lpq', Q0, Q0)))                  # F_pq against sqrt(F_pp F_qq): off-diagonals may cancel
    assert np.max(np.abs(Q - Q0) / scale) < RTOL
    assert np.max(np.abs(F - F0) / np.sqrt(np.outer(np.diag(F0), np.diag(F0)))) < RTOL
    assert np.array_equal(F, F.T)                                               # one triangle computed, mirrored
    assert np.array_equal(field_fisher(n_f, spectra, tot, dC, w), F)            # deterministic (fixed-order sums)


def test_field_form_equals_reference_spectra_form(ctx, golden_dir):
    """Fields (kappa, g), spectra (kk, kg, gg), w = nmodes (2L+1) fsky: the interloper.py:200-253 Fisher matrix written by the
    unmodified reference (golden) and by the device spectra-form path."""
    from lbgcmb_b200 import fisher
    g = np.load(os.path.join(golden_dir, 'fisher.npz'))
    L, nm, fsky = g['Llls'], g['nmodes'], float(g['fsky'])
    cut = L < float(g['LllMax'])                                                # interloper.py:221
    s8, b1 = float(g['fid_sig8']), float(g['fid_b1'])
    tot = np.array([g['C_kk'] + g['N_kk'], g['C_kg'], g['C_gg'] + g['N_gg']])
    z = np.zeros_like(g['C_kk'])
    dC = np.array([[2. / s8 * g['C_kk'], 2. / s8 * g['C_kg'], 2. / s8 * g['C_gg']], [z, g['C_kg'] / b1, 2. * g['C_gg'] / b1]])   # dp_DL, :65-74
    w = nm * (2. * L + 1.) * fsky * cut
    F = fisher.field_fisher(2, [(0, 0), (0, 1), (1, 1)], tot, dC, w)
    assert np.max(np.abs(F / g['Fisher'] - 1)) < RTOL


def test_field_fisher_errors(ctx):
    from lbgcmb_b200.fisher import field_fisher
    one = np.ones((1, 3))
    with pytest.raises(ValueError, match='auto spectrum'):
        field_fisher(2, [(0, 1)], one, one[None], np.ones(3))
    with pytest.raises(ValueError, match='listed twice'):
        field_fisher(2, [(0, 0), (0, 1), (1, 0)], np.ones((3, 3)), np.ones((1, 3, 3)), np.ones(3))
    with pytest.raises(ValueError, match='outside'):
        field_fisher(2, [(0, 0), (1, 2)], np.ones((2, 3)), np.ones((1, 2, 3)), np.ones(3))
    with pytest.raises(ValueError, match='n_par'):
        field_fisher(1, [(0, 0)], one, np.ones((9, 1, 3)), np.ones(3))
    with pytest.raises(ValueError):
        field_fisher(1, [(0, 0)], one, np.ones((1, 2, 3)), np.ones(3))
    with pytest.raises(np.linalg.LinAlgError):
        field_fisher(2, [(0, 0), (0, 1), (1, 1)], np.array([[1., 1.], [1., 1.], [1., 1.]]), np.ones((1, 3, 2)), np.ones(2))   # singular C_l
    assert np.array_equal(field_fisher(1, [(0, 0)], np.ones((1, 0)), np.ones((2, 1, 0)), np.ones(0)), np.zeros((2, 2)))


def test_amplitude_parameters_described_instead_of_tabulated(ctx):
    """lbg_field_fisher_amp: rows of amplitude parameters formed on the device from (factor matrix, signal spectra) == the same rows
    handed over explicitly (bit for bit: the product amp * sig is the same double either way), also with no dense parameter at all."""
    from lbgcmb_b200.fisher import field_fisher, field_fisher_amp
    rng = np.random.default_rng(7)
    spectra, tot, dC, w = _random_case(rng, 25, 37, 2, True)
    n_spec = len(spectra)
    sig = tot * rng.uniform(0.3, 0.9, tot.shape)
    amp = np.zeros((3, n_spec))
    amp[0, 1:9] = 1.0; amp[0, 25:33] = 2.0
    amp[1, 9:17] = 1.0; amp[1, 33:41] = 2.0
    amp[2, 17:25] = 0.7; amp[2, 41:49] = 1.4
    full = np.concatenate([dC, amp[:, :, None] * sig[None]])
    want = field_fisher(25, spectra, tot, full, w)
    got = field_fisher_amp(25, spectra, tot, dC, amp, sig, w)
    assert got.shape == (5, 5) and np.array_equal(got, want)
    only = field_fisher_amp(25, spectra, tot, None, amp, sig, w)
    assert np.array_equal(only, want[2:, 2:])
    with pytest.raises(ValueError):
        field_fisher_amp(25, spectra, tot, dC, amp[:, :-1], sig, w)
