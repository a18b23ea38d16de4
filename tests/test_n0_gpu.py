"""GPU parity: lbg_n0_ho02 (csrc/n0.cu) through the Python mirror of cmb/Nkk.py:31 vs
  (a) golden vectors produced by the UNMODIFIED reference (tests/golden/n0_root_*.npz, oracle/make_golden.py),
  (b) the numpy oracle (oracle/n0.py) on the same seeded inputs,
  (c) size-independent properties at BASELINE grid sizes.
Tolerance: relative 1e-10 (BASELINE.json north_star), FP64."""
import glob
import os

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
RTOL = 1e-10


def _interps(ell, lens, nolens):
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    return cl_interps_from_tables(ell, lens, nolens)


def test_spline_eval_matches_interp1d(ctx, cl_tables):
    from scipy.interpolate import interp1d
    from lbgcmb_b200.prep_camb import ClInterp
    ell, lens, _ = cl_tables
    x = np.random.default_rng(5).uniform(2, ell[-1], 20000)
    for k in range(3):
        got = ClInterp(ell, lens[k])(x)
        ref = interp1d(ell, lens[k], kind='cubic', bounds_error=False, fill_value=0.0)(x)
        assert np.max(np.abs(got / ref - 1)) < 1e-13
    f = ClInterp(ell, lens[0])
    assert f(np.array([-1.0, ell[-1] + 0.5])).tolist() == [0.0, 0.0]     # fill_value=0.0 outside the table
    assert isinstance(f(10.0), float)


@pytest.mark.parametrize('fold', [False, True])
def test_golden_reference_vectors(ctx, golden_dir, fold):
    from lbgcmb_b200 import Nkk
    files = sorted(glob.glob(os.path.join(golden_dir, 'n0_root_*.npz')))
    assert files, 'golden fixtures missing'
    for f in files:
        g = np.load(f)
        li, ni = _interps(g['ell'], g['lensed'], g['unlensed'])
        out, _ = Nkk.n0_terms(li, ni, g['Ls'], [str(t) for t in g['terms']], float(g['thetab']), float(g['DeltaT']),
                              bool(g['iterative']), fold_phi=fold, Tcmb=float(g['Tcmb']),
                              cos_sin=(g['cos_phi'], g['sin_phi']))
        rel = np.max(np.abs(out / g['nkk_reference'] - 1))
        assert rel < RTOL, '%s fold=%s rel=%.3e' % (os.path.basename(f), fold, rel)


def test_against_live_oracle(ctx, cl_tables):
    from oracle import n0 as on0
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    Ls = np.array([50, 517, 5000])
    ref = on0.nkk_root(on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens), Ls, ('TT', 'TE', 'EE', 'EB'),
                       1.4, 1.0, True)
    li, ni = _interps(ell, lens, nolens)
    for fold in (False, True):
        got = Nkk.Nkk(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, iterative=True, pickleit=False) if fold else \
            Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=False)[0]
        assert np.max(np.abs(got / ref - 1)) < RTOL


def test_single_estimators_and_bb_against_oracle(ctx, cl_tables):
    from oracle import n0 as on0
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    nolens = nolens.copy()
    nolens[2] = 0.5 * lens[2]          # non-zero unlensed BB so the BB estimator is exercised
    Ls = np.array([333])
    oi = (on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens))
    li, ni = _interps(ell, lens, nolens)
    for terms in (['TT'], ['TE'], ['EE'], ['EB'], ['BB']):
        ref = on0.nkk_root(oi[0], oi[1], Ls, tuple(terms), 3.0, 5.0, False)
        got = Nkk.n0_terms(li, ni, Ls, terms, 3.0, 5.0, False)[0]
        assert np.max(np.abs(got / ref - 1)) < RTOL, terms


def test_grid_A_properties(ctx, cl_tables):
    """Full reference grid (60 L, 50..5000): fold == unfold, min-variance sum of the per-estimator integrals."""
    from lbgcmb_b200 import Nkk
    from lbgcmb_b200.prep_Llls import prep_Llls
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    _, Llls, _ = prep_Llls(60, 50, 5000, log10=True)
    a, ta = Nkk.n0_terms(li, ni, Llls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=True)
    b, tb = Nkk.n0_terms(li, ni, Llls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=False)
    assert np.all(np.isfinite(a)) and np.all(a > 0)
    assert np.max(np.abs(a / b - 1)) < 1e-11
    tot = ta['TT'] + ta['TE'] + ta['EE'] + 2.5 * ta['EB']
    pre = (Llls * (Llls + 1.) / 2.) ** 2.
    assert np.max(np.abs(pre * (2 * np.pi) ** 2 / tot / a - 1)) < 1e-13
    # each estimator alone is noisier than the combination
    for t in ('TT', 'TE', 'EE', 'EB'):
        assert np.all(pre * (2 * np.pi) ** 2 / ta[t] > a)
    # deterministic: bitwise identical on a second run
    a2, _ = Nkk.n0_terms(li, ni, Llls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=True)
    assert np.array_equal(a, a2)
    # 4 special rows (l1 = 1, 1.5 exact-u rows; 3000, 5000 cut rows) run on all 201 phi nodes, the rest on 101 mirror pairs
    assert Nkk.last_stats['nodes'] == 60 * (101 * (9999 - 4) + 201 * 4)


def test_edge_cases(ctx, cl_tables):
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    assert Nkk.Nkk(li, ni, np.array([], dtype=int), pickleit=False).shape == (0,)
    with pytest.raises(KeyError):
        Nkk.Nkk(li, ni, np.array([100]), terms=['TB'], pickleit=False)
    with pytest.raises(ValueError):
        Nkk.Nkk(li, ni, np.array([100]), terms=['XY'], pickleit=False)
    one = Nkk.Nkk(li, ni, np.array([100]), pickleit=False)
    many = Nkk.Nkk(li, ni, np.array([100, 100, 4000, 100]), pickleit=False)
    assert many[0] == one[0] == many[1] == many[3]


def test_full_grid_B_properties(ctx, cl_tables):
    """BASELINE.json configs[1] at full size (every integer L in [50, 5000]): size-independent properties.
    * every L of the run equals the same L evaluated in a small call (the value of an L never depends on its neighbours);
    * the range-walking kernel (rows5, default) and the per-node-mask kernel (rows3, fold_phi=2) agree to the last bits on all 4951 L
      and count the same unmasked nodes;
    * N0 is smooth in L: no isolated outliers (a dropped or doubled node shows up as a spike at the 1e-3 level);
    * three spot values against the live numpy oracle."""
    from lbgcmb_b200 import Nkk
    from oracle import n0 as on0
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    Ls = np.arange(50, 5001)
    full, tf = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True)
    assert full.shape == (4951,) and np.all(np.isfinite(full)) and np.all(full > 0)
    pick = np.array([50, 51, 999, 1000, 2999, 3000, 3001, 4999, 5000])
    small, _ = Nkk.n0_terms(li, ni, pick, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True)
    assert np.array_equal(full[pick - 50], small)
    assert Nkk.last_stats['path'] == 5
    v3, t3 = Nkk.n0_terms(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True, fold_phi=2)      # generic kernel, same folded grid
    assert Nkk.last_stats['path'] == 3
    assert np.max(np.abs(full / v3 - 1)) < 5e-15
    for t in ('TT', 'TE', 'EE', 'EB'):
        assert np.max(np.abs(tf[t] / t3[t] - 1)) < 5e-15
    lf = np.log(full)
    second = lf[2:] - 2. * lf[1:-1] + lf[:-2]
    assert np.max(np.abs(second)) < 1e-3
    spot = np.array([137, 2222, 4321])
    ref = on0.nkk_root(on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens), spot, ('TT', 'TE', 'EE', 'EB'), 1.4, 1.0, True)
    assert np.max(np.abs(full[spot - 50] / ref - 1)) < 1e-10


def test_async_lane_matches_sync_and_overlaps_limber(ctx, cl_tables):
    """Nkk.n0_terms_async / Nkk_async (second context of the same GPU): bit-identical to the blocking call, also when Limber
    launches run on the default context while it is in flight."""
    from lbgcmb_b200 import Nkk, pmh
    from lbgcmb_b200.lensing import Ckk
    from lbgcmb_b200.synthetic import eh_like_plin
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    Ls = np.arange(50, 450)
    want, want_terms = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False)
    k, p = eh_like_plin()
    Pk = pmh.get_PkInterps(pmh.PkTable(k, p))
    ckk0 = Ckk(Pk, Ls.astype(np.float64))
    fut = Nkk.n0_terms_async(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False)
    ckk1 = [Ckk(Pk, Ls.astype(np.float64)) for _ in range(3)]          # default context, while N0 runs on lane 1
    got, got_terms = fut.result()
    assert np.array_equal(got, want)
    for t in want_terms:
        assert np.array_equal(got_terms[t], want_terms[t])
    for c in ckk1:
        assert np.array_equal(c, ckk0)
    assert np.array_equal(Nkk.Nkk_async(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 7.0, False).result(), want)


def test_kernel_paths_and_unmasked_node_count(ctx, cl_tables):
    """Which integrand kernel runs is decided by the inputs and reported (no environment switch): the range-walking kernel (5) for
    the folded reference grid, the generic one (3) for an unfolded grid, on request, for a table that ends in range and for a beam
    beyond the folded-noise bound.  The kernels count the nodes that pass a mask: equal to a direct count of the reference's masks
    (cmb/Nkk.py:13-26) on the same grid, for both kernels and for folded / unfolded grids."""
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    Ls = np.array([50, 333, 2999, 3001, 5000])
    want = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False)[0]
    assert Nkk.last_stats['path'] == 5
    n5 = Nkk.last_stats['nodes_unmasked']
    got = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False, fold_phi=2)[0]
    assert Nkk.last_stats['path'] == 3 and Nkk.last_stats['nodes_unmasked'] == n5
    assert np.max(np.abs(got / want - 1)) < 1e-12
    got = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False, fold_phi=False)[0]
    assert Nkk.last_stats['path'] == 3
    n_full = Nkk.last_stats['nodes_unmasked']
    assert np.max(np.abs(got / want - 1)) < 1e-11
    # direct count of the reference's masks on the full grid (P family: |u| <= 5000 always holds; 2 <= |v| <= 5000)
    l1, _, pphi, _ = Nkk.root_grids()
    cnt = 0
    for L in Ls:
        ux, uy = l1[:, None] * np.cos(pphi)[None, :], l1[:, None] * np.sin(pphi)[None, :]
        nv = ((L - ux) ** 2. + uy ** 2.) ** 0.5
        nu = (ux ** 2. + uy ** 2.) ** 0.5
        cnt += int(np.count_nonzero((nv >= 2) & (nv <= 5000) & (nu <= 5000)))
    assert n_full == cnt
    # a folded node stands for two nodes (one on the axis): 2 n5 - (nodes at phi = 0, pi) - (special rows counted on the full grid) ~ n_full
    assert abs(2 * n5 - n_full) < 0.02 * n_full
    Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 9.0, 30.0, False)
    assert Nkk.last_stats['path'] == 3          # beam beyond the folded-noise bound: exp() per node


def test_concurrent_lanes_with_different_grids(ctx, cl_tables):
    """Two N0 calls in flight on the two contexts of one GPU with DIFFERENT phi grids and kernels (lane 1: folded, range-walking;
    lane 0: unfolded, generic): each lane has its own block of the constant-memory phi table, so neither disturbs the other."""
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    Ls = np.arange(50, 850)
    a0 = Nkk.n0_terms(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False)[0]
    b0 = Nkk.n0_terms(li, ni, Ls[:60], ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False, fold_phi=False)[0]
    ca = Nkk.n0_begin(li, ni, Ls, ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False, lane=1)
    cb = Nkk.n0_begin(li, ni, Ls[:60], ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False, fold_phi=False, lane=0)
    with pytest.raises(Exception):
        Nkk.n0_begin(li, ni, Ls[:3], ('TT',), 1.4, 7.0, False, lane=0)          # one call in flight per context
    assert np.array_equal(cb.result()[0], b0)
    assert np.array_equal(ca.result()[0], a0)
    assert np.array_equal(Nkk.n0_terms(li, ni, Ls[:3], ('TT', 'TE', 'EE', 'EB'), 1.4, 7.0, False)[0], a0[:3])   # the context is usable again


def test_flat_spectrum_tt_closed_form(ctx):
    """SURVEY.md 8(c) closed form: constant C_l (lensed = unlensed = c0), white noise without a beam and no |v| cuts.  Then
    f_TT = c0 (L.u + L.v) = c0 L^2 for every node, F = f / (2 T^2) with T = c0 + N, the integrand is constant and the reference's
    Simpson rule over l1 in [1, 5000] x phi in [0, 2 pi] is exact for it:
        integral = c0^2 L^4 / (2 T^2) * pi (5000^2 - 1),   N0 = (L (L + 1) / 2)^2 (2 pi)^2 / integral.
    Non-grid L so that v = L - u never vanishes.  Independent of the oracle."""
    from lbgcmb_b200 import Nkk
    from lbgcmb_b200.DetectorNoise import DetectorNoise
    c0 = 3.0e-10
    n = 10151
    ell = np.arange(n, dtype=np.float64)
    tab = np.zeros((4, n))
    tab[0] = c0                                   # TT flat; EE, BB, TE zero
    li, ni = _interps(ell, tab, tab.copy())
    Ls = np.array([50.25, 777.3, 2999.9, 4999.7])
    DeltaT = 5.0
    noise = float(DetectorNoise(np.array([100.]), 0.0, DeltaT, 'TT')[0])       # thetab = 0: white
    T = c0 + noise
    want = (Ls * (Ls + 1.) / 2.) ** 2. * (2. * np.pi) ** 2. / (c0 ** 2. * Ls ** 4. / (2. * T ** 2.) * np.pi * (5000. ** 2. - 1.))
    got = Nkk.n0_terms(li, ni, Ls, ['TT'], 0.0, DeltaT, False, lmax_T=1e9, lmax_P=1e9, lmin_v=1e-9, exact_u_below=0.0)[0]
    assert np.max(np.abs(got / want - 1)) < 1e-12
    nofold = Nkk.n0_terms(li, ni, Ls, ['TT'], 0.0, DeltaT, False, lmax_T=1e9, lmax_P=1e9, lmin_v=1e-9, exact_u_below=0.0,
                          fold_phi=False)[0]
    assert np.max(np.abs(nofold / want - 1)) < 1e-12


def test_short_and_long_tables(ctx, cl_tables):
    """Tables that end inside the integration range (the reference's interp1d returns 0 beyond its last knot, so C^tot is pure
    noise there: prep_camb.py:88-89; generic kernel with the zero-extended table) and tables longer than the per-call table's
    compile-time stride (only the intervals |v| can reach are tabulated: still the range-walking kernel)."""
    from oracle import n0 as on0
    from lbgcmb_b200 import Nkk
    from lbgcmb_b200.synthetic import cmb_cl_tables
    ell, lens, nolens = cl_tables
    n = 4001                                                   # knots 0..4000 < lmax_P = 5000
    Ls = np.array([60, 3500])
    ref = on0.nkk_root(on0.cl_interps(ell[:n], lens[:, :n]), on0.cl_interps(ell[:n], nolens[:, :n]), Ls, ('TT', 'TE', 'EE', 'EB'),
                       1.4, 7.0, False)
    li, ni = _interps(ell[:n], lens[:, :n], nolens[:, :n])
    got = Nkk.Nkk(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 7.0, iterative=False, pickleit=False)
    assert np.max(np.abs(got / ref - 1)) < RTOL
    assert Nkk.last_stats['path'] == 3
    ell2, lens2, nolens2 = cmb_cl_tables(n_ell=17001)          # > 16382 knots
    assert np.array_equal(lens2[:, :ell.size], lens)           # same spectra, longer tables
    li2, ni2 = _interps(ell2, lens2, nolens2)
    li0, ni0 = _interps(ell, lens, nolens)
    Ls = np.array([50, 2000, 5000])
    a = Nkk.Nkk(li2, ni2, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, iterative=True, pickleit=False)
    assert Nkk.last_stats['path'] == 5
    b = Nkk.Nkk(li0, ni0, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, iterative=True, pickleit=False)
    assert np.max(np.abs(a / b - 1)) < RTOL                    # the far end of a not-a-knot spline does not reach l <= 10000


def test_unusual_L_and_wide_beam_against_oracle(ctx, cl_tables):
    """L outside the usual grid (below the first l1 node, non-integer, beyond lmax_P where only part of the |v| annulus survives)
    and the widest beam / highest noise of the BASELINE sweep (the noise factor reaches e^19 at l = 5000); a 10 arcmin beam leaves
    the folded-noise bound and runs the analytic-noise variant of the generic kernel."""
    from oracle import n0 as on0
    from lbgcmb_b200 import Nkk
    ell, lens, nolens = cl_tables
    li, ni = _interps(ell, lens, nolens)
    ol, on = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    terms = ('TT', 'TE', 'EE', 'EB')
    for Ls, thetab, DeltaT, it in ((np.array([2, 7.5, 6000, 9990]), 1.4, 7.0, False), (np.array([50, 1234, 5000]), 7.0, 30.0, False),
                                   (np.array([3000, 3001]), 1.0, 1.0, True), (np.array([50, 2500]), 10.0, 30.0, False)):
        ref = on0.nkk_root(ol, on, Ls, terms, thetab, DeltaT, it)
        got = Nkk.n0_terms(li, ni, Ls, terms, thetab, DeltaT, it)[0]
        assert np.all(np.isfinite(got)) and np.all(got > 0)
        assert np.max(np.abs(got / ref - 1)) < RTOL, (Ls, thetab, np.abs(got / ref - 1))
