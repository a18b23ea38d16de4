"""GPU parity: the cadence/cadence.py mirror (lbg_cadence_fisher + thin-shell Limber C_gg per shell) against the golden
vector written by the UNMODIFIED reference function (tests/golden/cadence.npz, oracle/make_golden.py:make_cadence).  Tolerance: relative 1e-10 on the shell counts and the Schur-Limber errors;
the Limber column goes through the inverse of the 6 x 6 Fisher matrix and is held to 1e-8."""
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_cadence_fisher_golden(ctx, golden_dir):
    from lbgcmb_b200 import cadence, pmh
    from oracle.make_golden import cadence_inputs
    g = np.load(os.path.join(golden_dir, 'cadence.npz'))
    Pk = pmh.get_PkInterps(pmh.PkTable(g['k'], g['plin']))
    T = cadence_inputs()
    assert np.allclose([float(cadence.lenses_bz(z)) for z in g['zs']], g['lenses_bz'], rtol=1e-13)
    assert np.max(np.abs(cadence.Cij(Pk, g['L'][:50].astype(np.float64), 0.45, 0.55) / g['Cij'] - 1)) < 1e-10
    cadence.dz = 0.1
    for name, sources in (('lenses', False), ('sources', True)):
        got = cadence.Fisher(Pk, g['L'], g['zs'], float(g['tNs']), float(g['tNp']), T['pz'], T['bz'], fsky=float(g['fsky']),
                             fover=float(g['fover']), sources=sources, printit=False)
        ref = g['out_' + name]
        assert got.shape == ref.shape == (g['zs'].size, 5)
        assert np.max(np.abs(got[:, :4] / ref[:, :4] - 1)) < 1e-10, name
        assert np.max(np.abs(got[:, 4] / ref[:, 4] - 1)) < 1e-8, name
    assert cadence.last['LMAX'] == int(np.ceil(0.3 * __import__('lbgcmb_b200.utils', fromlist=['x']).comoving_distance(g['zs'][-1]) - 0.5))
