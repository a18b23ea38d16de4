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


def test_growth_rate_closed_form_and_fsigma8_propagation():
    """f = d ln D / d ln a from the growth table's closed form against the numerical derivative of the same table (1e-6) and the
    Omega_m(a)^0.55 approximation (1 % in LCDM); with sigma8 the only cosmological parameter, f sigma8 is proportional to it and the
    propagated relative error equals sigma(sigma8) / sigma8."""
    from lbgcmb_b200 import tomography as tm
    c = tm.Background(M=0.292, B=0.0468, sig8=0.819)                   # host Romberg growth table (no device)
    z = np.array([0.1, 0.5, 1., 2., 4.])
    a = 1. / (1. + z)
    f = tm.growth_rate(c, z)
    h = 1e-4
    num = (np.log(c.Dplus(a * np.exp(h))) - np.log(c.Dplus(a * np.exp(-h)))) / (2. * h)
    assert np.max(np.abs(num / f - 1)) < 1e-6
    assert np.max(np.abs(f / c.OmegaMa(z=z) ** 0.55 - 1)) < 1e-2
    assert np.allclose(tm.fsigma8(c, z), f * 0.819 * c.Dgrow(z=z), rtol=1e-15)
    result = {'params': ['sig8', 'b_x'], 'fiducial': dict(tm.FIDUCIAL), 'cov': np.array([[4e-6, 1e-7], [1e-7, 9e-6]])}
    r = tm.fsigma8_forecast(result, z, growth='host')
    assert np.max(np.abs(r['sigma'] / r['fsigma8'] - 2e-3 / 0.819)) < 1e-12 and np.all(r['jacobian'][:, 1] == 0.0)


# ---- sharding logic of tomographic_fisher over gloo (world sizes 2 and 3); the device calls are replaced by numpy stand-ins ----------
import os
import sys

import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
TRACERS = [{'name': 'a', 'edges': [0.5, 0.7, 0.9], 'nbar': [2e6, 3e6]}, {'name': 'b', 'edges': [1.0, 1.5], 'nbar': 1e6}]
ELL = np.arange(50., 91.)


def _toy_spectra(theta, ell):
    """Smooth stand-in spectra with a parameter dependence: kk, kg [3], gg [3]."""
    s8, mnu = theta['sig8'], theta['mnu']
    shape = (ell / 100.) ** -1.2 * (1. - 0.3 * mnu * np.log(ell / 50. + 1.))
    ckk = 1e-7 * s8 ** 2. * shape
    ckg = np.array([2e-7, 3e-7, 1.5e-7])[:, None] * s8 ** 2. * shape
    cgg = np.array([2e-6, 3e-6, 4e-6])[:, None] * s8 ** 2. * shape * (1. + 0.1 * mnu)
    return ckk, ckg, cgg


def _dense_accum(nf, spectra, tot, dC, w):
    from oracle.fisher import field_fisher
    return field_fisher(nf, spectra, tot, dC, w)[0] if len(w) else np.zeros((dC.shape[0], dC.shape[0]))


def _run(sharded):
    from lbgcmb_b200 import tomography as tm
    return tm.tomographic_fisher(TRACERS, ELL, np.ones_like(ELL), 0.4, 3e-8 * np.ones_like(ELL), cosmo_params=('sig8', 'mnu'),
                                 sharded=sharded, spectra_fn=_toy_spectra, accum=_dense_accum)


def _worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port))
    dist.init_process_group('gloo', rank=rank, world_size=world)
    r = _run(True)
    q.put((rank, r['F'], r['spectra']['kk'].size))
    dist.barrier()
    dist.destroy_process_group()


def _spawn(world, port):
    ctx = mp.get_context('spawn')
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    out = sorted(q.get(timeout=120) for _ in range(world))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    return out


def test_sharded_tomographic_fisher_gloo():
    ref = _run(False)
    assert ref['params'] == ['sig8', 'mnu', 'b_a', 'b_b'] and ref['fields'] == 4
    assert np.all(np.linalg.eigvalsh(ref['F']) > 0)
    for world, port in ((2, 29611), (3, 29612)):
        out = _spawn(world, port)
        assert sum(o[2] for o in out) == ELL.size                           # the ranks' multipoles partition the grid
        for rank, F, n in out:
            assert np.max(np.abs(F / ref['F'] - 1)) < 1e-12                 # same matrix on every rank
