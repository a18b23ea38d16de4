"""Tomographic workload of BASELINE.json configs[3] at its stated size: 6 tracers x 20 redshift bins (dz = 0.1) on 4951 multipoles.

  python scripts/tomography_bench.py            (one GPU; prints one JSON line)

The reference has no single driver for it: 6 tracers = yu.py:44-48, 20 bins of dz = 0.1 = mqw.py:20, the
(sigma8, b1) Fisher = interloper.py:200-260.  Here: all 120 bin samples' C_gg and C_kg from TWO batched Limber launches
(sweep.limber_grid), C_kk, one N0 curve, then S/N + Fisher for the 120 bins in ONE launch (sweep.forecast_grid); three randomly
chosen bins are recomputed through the per-sample mirrors (Cgg.Cgg / Ckg.Ckg / snr.snr / interloper.fisher_bias(fused=False))
and must agree to 1e-11.  Tracers are synthetic stand-ins (Gaussian n(z), b(z) = b0 (1 + z) / (1 + z0)) for DESI BGS / LRG / ELG /
QSO and two LSST LBG samples; the reference's dat/ss17_tracers.dat is not available on the GPU box."""
import json
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

TRACERS = {   # name: (z0, sigma, nbar [deg^-2], b0)
    'BGS': (0.3, 0.15, 1400., 1.34), 'LRG': (0.75, 0.15, 600., 2.1), 'ELG': (1.0, 0.25, 2400., 1.2),
    'QSO': (1.7, 0.6, 170., 2.3), 'LBG-u': (2.96, 0.24, 2181., 3.96), 'LBG-g': (3.8, 0.3, 5409., 4.8)}
N_BIN, DZ = 20, 0.1


def main():
    import bench
    from lbgcmb_b200 import Nkk, pmh, sweep, interloper, snr as snr_mod, _ffi
    from lbgcmb_b200.Cgg import Cgg, Ngg
    from lbgcmb_b200.Ckg import Ckg
    from lbgcmb_b200.lensing import Ckk
    from lbgcmb_b200.prep_camb import cl_interps_from_tables
    from lbgcmb_b200.synthetic import gaussian_pz
    li, ni = cl_interps_from_tables(*bench.make_inputs())
    kk, pp, _ = bench.forecast_inputs()
    Pk = pmh.get_PkInterps(pmh.PkTable(kk, pp))
    Ls = np.arange(50, 5001)
    Lf = Ls.astype(np.float64)
    nmodes = np.ones_like(Lf)
    samples = []
    for name, (z0, sig, nbar, b0) in TRACERS.items():
        full = gaussian_pz(z0, sig)
        lo0 = max(z0 - N_BIN * DZ / 2., 0.02)
        norm = np.trapezoid(full(np.arange(0.01, 10.0, 0.001)), dx=0.001)
        for i in range(N_BIN):
            zlo, zhi = lo0 + i * DZ, lo0 + (i + 1) * DZ
            pz = (lambda f, a, b: (lambda z: np.where((np.asarray(z) >= a) & (np.asarray(z) < b), f(z), 0.0)))(full, zlo, zhi)
            frac = np.trapezoid(full(np.arange(zlo, zhi, 0.001)), dx=0.001) / norm
            bz = (lambda b_, z_: (lambda z: b_ * (1. + z) / (1. + z_)))(b0, z0)
            samples.append({'name': '%s[%d]' % (name, i), 'pz': pz, 'bz': bz, 'nbar': max(nbar * frac, 1e-3),
                            'fid_b1': float(bz(0.5 * (zlo + zhi))), 'LllMax': 2000.0})
    ctx = _ffi.context()
    nkk = Nkk.Nkk(li, ni, Ls, ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, iterative=True, pickleit=False)
    sweep.limber_grid(Pk, Lf[:16], samples[:9])                        # warm-up (allocations)
    ctx.kernel_totals(reset=True)
    t0 = time.perf_counter()
    ckk = Ckk(Pk, Lf)
    cgg, ckg, ngg = sweep.limber_grid(Pk, Lf, samples)
    t1 = time.perf_counter()
    out = sweep.forecast_grid(Lf, nmodes, ckk, cgg, ckg, ngg, nkk[None, :], bench.FSKY, 0.819, [s['fid_b1'] for s in samples],
                              [s['LllMax'] for s in samples])
    t2 = time.perf_counter()
    kms, launches = ctx.kernel_totals()
    assert np.all(np.isfinite(out)) and np.all(out[0, :, 0] >= 0)
    worst = 0.0
    for i in np.random.default_rng(3).choice(len(samples), 3, replace=False):
        s = samples[i]
        a = Cgg(Pk, Lf, 0.01, 10.0, s['pz'], s['bz'], zeff=False)
        b = Ckg(Pk, Lf, 0.01, 10.0, s['pz'], s['bz'], zeff=False)
        n = Ngg(Lf, 0.01, 10.0, s['pz'], s['nbar'])
        Cls = {'gg': a, 'kg': b, 'gk': b, 'kk': ckk}
        Nls = {'gg': n, 'kg': 0 * b, 'gk': 0 * b, 'kk': nkk}
        fb = interloper.fisher_bias(Lf, nmodes, Cls, Nls, Cls, Nls, bench.FSKY, 0.819, s['fid_b1'], s['LllMax'], fused=False)
        vkg = ((ckk + nkk) * (a + n) + b ** 2.) / ((2. * Lf + 1.) * bench.FSKY)
        sn = snr_mod.snr(Lf, b, vkg, nmodes, lmax=s['LllMax'])
        worst = max(worst, np.max(np.abs(cgg[i] / a - 1)), np.max(np.abs(ckg[i] / b - 1)), abs(ngg[i] / n[0] - 1),
                    np.max(np.abs(out[0, i, 1:5].reshape(2, 2) / fb['Fisher'] - 1)), abs(out[0, i, 0] / sn - 1))
    assert worst < 1e-11, worst
    n_s = len(samples)
    sn = out[0, :, 0].reshape(len(TRACERS), N_BIN)
    print(json.dumps({'workload': '%d tracers x %d bins (dz = %.1f) on %d multipoles: C_gg + C_kg (1000 / 999 z nodes), C_kk, S/N + (sigma8, b1) Fisher per bin'
                                  % (len(TRACERS), N_BIN, DZ, Ls.size),
                      'limber_wall_ms': (t1 - t0) * 1e3, 'fisher_wall_ms': (t2 - t1) * 1e3, 'kernel_ms': kms, 'launches': launches,
                      'limber_nodes': int(n_s * Ls.size * (1000 + 999) + Ls.size * 2000),
                      'limber_nodes_per_s_wall': (n_s * Ls.size * 1999 + Ls.size * 2000) / (t1 - t0),
                      'nodes_per_s_kernels': (n_s * Ls.size * 1999 + Ls.size * 2000) / (kms * 1e-3),
                      'max_rel_dev_vs_per_sample_mirrors': float(worst),
                      'best_bin_snr_per_tracer': {k: [int(np.argmax(sn[i])), float(sn[i].max())] for i, k in enumerate(TRACERS)}}), flush=True)


if __name__ == '__main__':
    main()
