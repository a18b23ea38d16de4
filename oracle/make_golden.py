"""Generate tests/golden/*.npz by running the UNMODIFIED reference (build container only).

TEST INFRASTRUCTURE.  usage:  python -m oracle.make_golden [n0|limber|halofit|modi|fisher ...]
Each fixture stores the exact inputs (tables, grids, cos/sin of the phi nodes as computed by this
machine's numpy) next to the reference's outputs, so the replay on the GPU box needs neither
/root/reference nor bit-identical libm.
"""
import os
import sys
import time

import numpy as np

from . import refload, n0 as on0

GOLD = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')


def make_n0():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    Nkk = refload.load('Nkk')                      # /root/reference/cmb/Nkk.py
    ell, lens, nolens = cmb_cl_tables(5300)
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    _, pphi, _, _ = on0.root_grids()
    cases = [
        ('n0_root_cmbs4', np.array([50, 3000]), ['TT', 'TE', 'EE', 'EB'], 1.4, 1.0, True),
        ('n0_root_so_ttee', np.array([1000]), ['TT', 'EE'], 1.4, 7.0, False),
        ('n0_root_planck_eb', np.array([4999]), ['EB', 'TE'], 7.0, 30.0, False),
    ]
    for name, Ls, terms, thetab, DeltaT, it in cases:
        t = time.time()
        ref = Nkk.Nkk(li, ni, Ls, terms, thetab, DeltaT, iterative=it, pickleit=False)
        dt = time.time() - t
        ora = on0.nkk_root(li, ni, Ls, tuple(terms), thetab, DeltaT, it)
        print(name, 'reference %.1f s' % dt, ref, 'oracle rel diff', np.max(np.abs(ora / ref - 1)))
        np.savez_compressed(os.path.join(GOLD, name + '.npz'), ell=ell, lensed=lens, unlensed=nolens, Ls=Ls,
                            terms=np.array(terms), thetab=thetab, DeltaT=DeltaT, iterative=it, Tcmb=2.786,
                            cos_phi=np.cos(pphi), sin_phi=np.sin(pphi), nkk_reference=ref, ref_seconds=dt)


def _pk_setup():
    from lbgcmb_b200.synthetic import eh_like_plin, PARAMS
    from lbgcmb_b200.modi.cosmology import Cosmology      # host-side growth D(z): an INPUT shared by GPU path and oracle
    from .shim.astropy.cosmology import FlatLambdaCDM
    k, p = eh_like_plin()
    eng = Cosmology(M=PARAMS['om_m'], H0=100., B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=k, plin=p)
    cosmo = FlatLambdaCDM(H0=100., Om0=PARAMS['om_m'], Ob0=PARAMS['om_b'])     # modi/cosmology.py:32
    return k, p, eng, cosmo, PARAMS


def make_halofit():
    from . import halofit as ohf
    hf = refload.load('halofit')                   # /root/reference/modi/halofit.py
    k, p, eng, cosmo, P = _pk_setup()
    zs = np.array([0.0, 0.5, 1.0, 2.0, 3.0, 4.0])
    D = np.array([float(eng.Dgrow(z=z)) for z in zs])
    out, spec, out_smith = [], [], []
    for z, d in zip(zs, D):
        delta = p * k ** 3 / (2. * np.pi ** 2) * d ** 2.
        out.append(hf.halofit(k, delta, P['sig_8'], z, cosmo, takahashi=True))
        out_smith.append(hf.halofit(k, delta, P['sig_8'], z, cosmo, takahashi=False))
        spec.append(hf._get_spec(k, delta, P['sig_8']))
        mine = ohf.halofit(k, delta, P['sig_8'], z, cosmo, True)
        print('halofit z=%.1f spec=%s oracle rel diff %.2e' % (z, spec[-1], np.max(np.abs(mine / out[-1] - 1))))
    np.savez_compressed(os.path.join(GOLD, 'halofit.npz'), k=k, plin=p, zs=zs, D=D, delta_nl=np.array(out),
                        delta_nl_smith=np.array(out_smith), spec=np.array(spec), sig8=P['sig_8'], om_m=P['om_m'], om_b=P['om_b'])


def make_limber():
    from . import halofit as ohf, limber as olim
    hf = refload.load('halofit')
    Cgg, Ckg, lensing = refload.load('Cgg'), refload.load('Ckg'), refload.load('lensing')   # reference root modules
    Gp = refload.load('Gaussian_pz')
    prep = refload.load('prep_Llls')
    k, p, eng, cosmo, P = _pk_setup()
    Dg = lambda z: float(eng.Dgrow(z=z))
    ref_pk = ohf.ModiPk(k, p, Dg, cosmo, P['sig_8'], halofit_fn=hf.halofit).interps()      # reference halofit inside
    ora_pk = ohf.ModiPk(k, p, Dg, cosmo, P['sig_8']).interps()
    bg = olim.Background(P)
    _, Llls, nmodes = prep.prep_Llls(60, 50, 5000, log10=True)
    pz_u = lambda z: Gp.Gaussian_pz(z, 2.96, 0.24)
    pz_g = lambda z: Gp.Gaussian_pz(z, 3.8, 0.3)
    bz = lambda z: 1. + z
    zmin, zmax = 0.96, 4.96
    res = {}
    t = time.time()
    res['cgg_zeff'] = Cgg.Cgg(ref_pk, Llls, zmin, zmax, pz_u, bz, zeff=True)
    res['cgg_full'] = Cgg.Cgg(ref_pk, Llls, zmin, zmax, pz_u, bz, zeff=False)
    res['cgg_ug_full'] = Cgg.Cgg(ref_pk, Llls, 1.8, 5.8, pz_u, bz, survey_pz2=pz_g, bz2=bz, zeff=False)
    res['ckg_zeff'] = Ckg.Ckg(ref_pk, Llls, zmin, zmax, pz_u, bz, zeff=True)
    res['ckg_full'] = Ckg.Ckg(ref_pk, Llls, zmin, zmax, pz_u, bz, zeff=False)
    res['ckk'] = lensing.Ckk(ref_pk, Llls)
    res['ngg'] = Cgg.Ngg(Llls, zmin, zmax, pz_u, 2181.0)
    res['lensing_kernel'] = lensing.lensing_kernel(np.array([0.5, 1.0, 3.0, 10.0]))
    print('reference limber calls: %.1f s' % (time.time() - t))
    chk = {'cgg_zeff': olim.Cgg(bg, ora_pk, Llls, zmin, zmax, pz_u, bz, zeff=True),
           'cgg_full': olim.Cgg(bg, ora_pk, Llls, zmin, zmax, pz_u, bz, zeff=False),
           'cgg_ug_full': olim.Cgg(bg, ora_pk, Llls, 1.8, 5.8, pz_u, bz, pz_g, bz, zeff=False),
           'ckg_zeff': olim.Ckg(bg, ora_pk, Llls, zmin, zmax, pz_u, bz, zeff=True),
           'ckg_full': olim.Ckg(bg, ora_pk, Llls, zmin, zmax, pz_u, bz, zeff=False),
           'ckk': olim.Ckk(bg, ora_pk, Llls), 'ngg': olim.Ngg(Llls, zmin, zmax, pz_u, 2181.0),
           'lensing_kernel': bg.lensing_kernel(np.array([0.5, 1.0, 3.0, 10.0]))}
    for name in res:
        print(name, 'oracle rel diff %.2e' % np.max(np.abs(chk[name] / res[name] - 1)))
    zchk = np.array([0.0, 0.96, 2.0, 2.96, 4.0, 4.96, 10.0, 1100.0])
    np.savez_compressed(os.path.join(GOLD, 'limber_root.npz'), k=k, plin=p, Llls=Llls, nmodes=nmodes, zmin=zmin, zmax=zmax,
                        zchk=zchk, Dchk=np.array([Dg(z) for z in zchk]), chichk=bg.chi(zchk), **res)


def fisher_tracers():
    """Analytic stand-ins for the survey inputs of interloper.py / mqw.py / yu.py (shared by the GPU tests)."""
    g = lambda z0, s: (lambda z: np.exp(-(np.asarray(z, dtype=np.float64) - z0) ** 2 / (2. * s * s)) / (np.sqrt(2. * np.pi) * s))
    return {
        'pz_u': g(2.96, 0.24), 'bz': (lambda z: 1. + z),
        'pz_int': (lambda z: 0.94 * g(2.96, 0.24)(z) + 0.06 * g(1.5, 0.2)(z)), 'bz_int': (lambda z: 0.97 * (1. + z)),
        'yu': [((lambda z: 1.3 + 0. * z), g(0.4, 0.15), 900.), ((lambda z: 1.7 + 0. * z), g(0.8, 0.2), 300.),
               ((lambda z: 1.2 + 0.5 * z), g(1.4, 0.3), 150.)],
        'mqw_ps': g(3.0, 0.4), 'mqw_pp': g(3.1, 0.45), 'mqw_bs': (lambda z: 2.0 + 0.3 * z), 'mqw_bp': (lambda z: 1. + z),
    }


def make_fisher():
    import os as _os
    from . import halofit as ohf, fisher as ofi
    hf = refload.load('halofit')
    Cgg, Ckg, lensing = refload.load('Cgg'), refload.load('Ckg'), refload.load('lensing')
    snr = refload.load('snr')
    prep = refload.load('prep_Llls')
    sys.path[:0] = [_os.path.join(refload.REF, x) for x in ('dropouts', 'mqw/py', 'dropouts/bz/py', 'dropouts/schechter')]
    cwd = _os.getcwd()
    _os.chdir(refload.REF)                         # schmittfull_nz.py reads dat/ss17_tracers.dat relative to the cwd
    interloper, mqw = refload.load('interloper'), refload.load('mqw')
    _os.chdir(cwd)
    k, p, eng, cosmo, P = _pk_setup()
    Dg = lambda z: float(eng.Dgrow(z=z))
    ref_pk = ohf.ModiPk(k, p, Dg, cosmo, P['sig_8'], halofit_fn=hf.halofit).interps()
    T = fisher_tracers()
    _, Llls, nmodes = prep.prep_Llls(60, 50, 5000, log10=True)
    zmin, zmax, fsky = 0.96, 4.96, 0.4
    out = {'k': k, 'plin': p, 'Llls': Llls, 'nmodes': nmodes, 'zmin': zmin, 'zmax': zmax, 'fsky': fsky}
    t = time.time()
    # ---- interloper.py: fiducial and contaminated spectra with the reference's own Limber code --------------------------
    nkk = 2.0e-8 * (1. + (Llls / 1500.) ** 2.)
    Ld = Llls.astype(np.float64)
    def allcls(pz, bz, nbar):
        cgg = Cgg.Cgg(ref_pk, Llls, zmin, zmax, pz, bz, zeff=False, bz2=bz, survey_pz2=pz)
        ngg = Cgg.Ngg(Ld, zmin, zmax, pz, nbar)
        ckg = Ckg.Ckg(ref_pk, Llls, zmin, zmax, pz, bz, zeff=False)
        ckk = lensing.Ckk(ref_pk, Llls)
        return ({'gg': cgg, 'kg': ckg, 'gk': ckg, 'kk': ckk},
                {'gg': ngg, 'kg': np.zeros_like(ckg), 'gk': np.zeros_like(ckg), 'kk': nkk})
    Cls, Nls = allcls(T['pz_u'], T['bz'], 2181.0)
    xCls, xNls = allcls(T['pz_int'], T['bz_int'], 2181.0 / 0.94)
    print('interloper spectra: %.1f s' % (time.time() - t))
    covs = interloper.get_ClsCov(Llls, Cls, Nls, fsky)
    fid_sig8, fid_b1, LllMax = 0.819 * Dg(2.96), 3.96, 1752.6
    fb = ofi.fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax,
                         get_ClsCov=interloper.get_ClsCov, dp_DL=interloper.dp_DL)      # script body over the reference functions
    mine = ofi.fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax)
    print('fisher_bias oracle rel diff', np.max(np.abs(mine['Fisher'] / fb['Fisher'] - 1)), np.max(np.abs(mine['dtheta'] / fb['dtheta'] - 1)))
    for name, d in (('C', Cls), ('N', Nls), ('xC', xCls), ('xN', xNls)):
        for key in ('gg', 'kg', 'kk'):
            out['%s_%s' % (name, key)] = d[key]
    keys = list(Cls)
    out['cov_keys'] = np.array([a + b for a in keys for b in keys])
    out['cov'] = np.array([covs[a + b] for a in keys for b in keys])
    out.update(fid_sig8=fid_sig8, fid_b1=fid_b1, LllMax=LllMax, Fisher=fb['Fisher'], iFisher=fb['iFisher'], interim=fb['interim'],
               dtheta=fb['dtheta'])
    # ---- snr.py ---------------------------------------------------------------------------------------------------------
    vkg = ((Cls['kk'] + nkk) * (Cls['gg'] + Nls['gg']) + Cls['kg'] ** 2.) / ((2. * Llls + 1.) * fsky)
    out['vkg'] = vkg
    out['snr_all'] = snr.snr(Llls, Cls['kg'], vkg, nmodes)
    out['snr_cut'] = snr.snr(Llls, Cls['kg'], vkg, nmodes, lmax=LllMax)
    # ---- mqw.py: 20 shells, every integer l in [60, 5000] -----------------------------------------------------------------
    _, Lm, _nm = prep.prep_Llls(60, 60, 5000, log10=False)
    t = time.time()
    zs = np.arange(2.0 + 0.05, 4.0 + 0.05, 0.1)
    res = mqw.Fisher(ref_pk, Lm, 50., 2000., T['mqw_ps'], T['mqw_bs'], T['mqw_pp'], T['mqw_bp'], dz=0.1, fsky=0.35, zmin=2.0, zmax=4.0,
                     fover=0.1, percentiles=list(zs))
    print('mqw.Fisher reference: %.1f s' % (time.time() - t))
    out['mqw_L'] = Lm
    out['mqw_out'] = np.array([float(np.ravel(x)[0]) for x in res])
    ub = lambda z: 1.0
    out['mqw_cgg'] = Cgg.Cgg(ref_pk, Lm, 2.0, 4.0, T['mqw_pp'], ub, survey_pz2=T['mqw_ps'], bz2=ub, zeff=True)
    # ---- yu.py: 3 tracers, 4 multipoles, z grid 0.01..10 --------------------------------------------------------------------
    t = time.time()
    Ly = Llls[::18]
    ckk_y = lensing.Ckk(ref_pk, Ly)
    rho, kgs, ggs, ngs = [], [], [], []
    for ii, L in enumerate(Ly):
        L1 = np.array([L])
        kg = [Ckg.Ckg(ref_pk, L1, 0.01, 10.0, pz, bz, zeff=False)[0] for (bz, pz, n) in T['yu']]
        gg = [[Cgg.Cgg(ref_pk, L1, 0.01, 10.0, pz, bz, bz2=b2, survey_pz2=p2, zeff=False)[0] for (bz, pz, n) in T['yu']] for (b2, p2, n2) in T['yu']]
        ng = [Cgg.Ngg(L1.astype(np.float64), 0.01, 10.0, pz, n)[0] for (bz, pz, n) in T['yu']]
        rho.append(ofi.rho_from_spectra(ckk_y[ii], kg, gg, ng))
        kgs.append(kg); ggs.append(gg); ngs.append(ng)
    print('yu: %.1f s' % (time.time() - t), rho)
    out.update(yu_L=Ly, yu_ckk=ckk_y, yu_kg=np.array(kgs), yu_gg=np.array(ggs), yu_ngg=np.array(ngs), yu_rho=np.array(rho))
    np.savez_compressed(os.path.join(GOLD, 'fisher.npz'), **out)


def make_modi():
    """modi/nkk.py run unmodified: N_fisher / N_fisherTE on in-memory splines, and the CMB class on CLASS-format files."""
    import tempfile
    from scipy.interpolate import InterpolatedUnivariateSpline as IUS
    from lbgcmb_b200.synthetic import cmb_cl_tables, write_class_files
    nkk = refload.load('nkk')                      # /root/reference/modi/nkk.py
    ell, lens, nolens = cmb_cl_tables(5801)
    eu, el = ell[2:5801], ell[2:5301]
    u = {'TT': nolens[0, 2:5801], 'EE': nolens[1, 2:5801], 'BB': nolens[2, 2:5801], 'TE': nolens[3, 2:5801]}
    le = {'TT': lens[0, 2:5301], 'EE': lens[1, 2:5301], 'BB': lens[2, 2:5301], 'TE': lens[3, 2:5301]}
    thetab, delT = 1.4, 7.0
    Ls = np.array([100., 1000., 10. ** 3.5])
    out = {'Ls': Ls, 'thetab': thetab, 'delT': delT}
    t = time.time()
    for mode in ('TT', 'EE', 'EB'):
        a1, a2 = mode[0] * 2, mode[1] * 2
        lint = np.linspace(1., 5000. if mode in ('EB', 'EE') else 3000., 10000)
        n1, n2 = on0.detector_noise(el, thetab, delT, a1, 2.786), on0.detector_noise(el, thetab, delT, a2, 2.786)
        args = (IUS(eu, u[a1]), IUS(eu, u[a2]), IUS(el, le[a1] + n1), IUS(el, le[a2] + n2))
        out['int_' + mode] = np.array([nkk.N_fisher(L, lint, *args, mode) for L in Ls])
        mine = np.array([on0.modi_N_fisher(L, lint, *args, mode) for L in Ls])
        print(mode, out['int_' + mode], 'oracle rel diff %.2e' % np.max(np.abs(mine / out['int_' + mode] - 1)))
    lint = np.linspace(1., 3000., 10000)
    n1, n2 = on0.detector_noise(el, thetab, delT, 'TT', 2.786), on0.detector_noise(el, thetab, delT, 'EE', 2.786)
    args = (IUS(eu, u['TE']), IUS(el, le['TT'] + n1), IUS(el, le['EE'] + n2), IUS(el, le['TE']))
    out['int_TE'] = np.array([nkk.N_fisherTE(L, lint, *args) for L in Ls])
    mine = np.array([on0.modi_N_fisherTE(L, lint, *args) for L in Ls])
    print('TE', out['int_TE'], 'oracle rel diff %.2e' % np.max(np.abs(mine / out['int_TE'] - 1)))
    print('N_fisher cases: %.1f s' % (time.time() - t))
    # the CMB container on files (modi/nkk.py:148-261), default 15-L grid
    t = time.time()
    d = tempfile.mkdtemp(prefix='lbgcmb_class_')
    clf, clsf = os.path.join(d, 'cl.dat'), os.path.join(d, 'cl_lensed.dat')
    write_class_files(clf, clsf)
    cmb = nkk.CMB(clfile=clf, clsfile=clsf)
    out['lat'] = cmb.lat
    out['cmb_TT'] = cmb.do_noise('T', 'T', thetab, delT)[2]
    out['cmb_EB'] = cmb.do_noise('E', 'B', thetab, delT)[2]
    out['cmb_TE'] = cmb.do_noiseTE(thetab, delT)[2]
    print('CMB class: %.1f s' % (time.time() - t), out['cmb_TT'][:3], out['cmb_EB'][:3], out['cmb_TE'][:3])
    np.savez_compressed(os.path.join(GOLD, 'n0_modi.npz'), **out)


def make_modi_lensing():
    """modi/cosmology.py + modi/lensing.py run unmodified (refload.load_modi): LensingX at z = 1 with an LSST-like dN/dz."""
    from . import modi_lensing as oml
    from lbgcmb_b200.synthetic import eh_like_plin, PARAMS
    cos, lens = refload.load_modi()
    k, p = eh_like_plin()
    c = cos.Cosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=k, plin=p)
    oc = oml.ModiCosmology(M=PARAMS['om_m'], B=PARAMS['om_b'], sig8=PARAMS['sig_8'], klin=k, plin=p)
    dndz = lambda z: z ** 2. * np.exp(-(z / 0.5) ** 1.0)
    dndz2 = lambda z: z ** 1.5 * np.exp(-(z / 0.4) ** 1.2)
    bz = lambda z: 1.438 + 0. * z                              # tools.BiasZ(cosmo).const(b = 1.438), modi/lensing.py:212
    ell = np.logspace(1, 3.47, 120)
    out = {'k': k, 'plin': p, 'l': ell, 'z': 1.0}
    res = {}
    for name, C, LX in (('ref', c, lens.LensingX), ('ora', oc, oml.LensingX)):
        t = time.time()
        mmz = lambda z: [k, C.pkanlin(z=z)[1]]
        hmz = lambda z: [k, C.pkanlin(z=z)[1] * bz(z)]
        hhz = lambda z: [k, C.pkanlin(z=z)[1] * bz(z) ** 2.]
        lx = LX(z=1.0, dndz=dndz, cosmo=C, l=ell)
        r = {'clkkz': lx.cl_kkz(mmz), 'clggz': lx.clz(hhz, auto=True), 'clkgz': lx.clz(hmz, auto=False),
             'clggzeff': lx.cl_zeff(hhz(1.0), auto=True), 'clkgzeff': lx.cl_zeff(hmz(1.0), auto=False),
             'clbitracer': lx.cl_zeff_bitracer(hhz(1.0), dndz2),
             'kerg': lx.kerg, 'kercmb': lx.kercmb, 'dggwtsdN': lx.dggwtsdN, 'dkgwtsdN': lx.dkgwtsdN, 'kkwts': lx.kkwts,
             'chis': lx.chis, 'norm': lx.norm}
        res[name] = r
        print(name, 'LensingX: %.1f s' % (time.time() - t))
    for key in res['ref']:
        print(key, 'oracle rel diff %.2e' % np.max(np.abs(np.asarray(res['ora'][key]) / np.asarray(res['ref'][key]) - 1)))
    zchk = np.array([0.0, 0.5, 1.0, 2.0, 4.0, 10.0, 1100.0])
    out.update(res['ref'])
    out.update(zchk=zchk, Dchk=np.array([c.Dgrow(z=z) for z in zchk]), chichk=np.array([c.chia(z=z) for z in zchk]),
               pknl_z1=c.pkanlin(z=1.0)[1], pklin_z5=c.pkanlin(z=5.0)[1])
    np.savez_compressed(os.path.join(GOLD, 'modi_lensing.npz'), **out)


def load_cadence():
    """The reference's cadence/cadence.py (functions only) loaded from where it lies: its two python-2 print statements (:174,
    :189) are parenthesised IN MEMORY and the __main__ block is not compiled; `np.array` on the ragged output rows (:186-187
    mix floats with 1-element arrays, an error on numpy >= 1.24) is given dtype=object through a proxy of the numpy module."""
    import re
    import types
    refload.setup()
    path = os.path.join(refload.REF, 'cadence', 'cadence.py')
    src = open(path).read().split("if __name__ == '__main__':")[0]
    src = re.sub(r'^(\s*)print  (.+)$', lambda m: '%sprint(%s)' % (m.group(1), m.group(2)), src, flags=re.M)
    mod = types.ModuleType('cadence_ref')
    mod.__file__ = path
    exec(compile(src, path, 'exec'), mod.__dict__)

    class _NP(object):
        def __getattr__(self, name):
            return getattr(np, name)

        @staticmethod
        def array(x, *a, **k):
            try:
                return np.array(x, *a, **k)
            except ValueError:
                return np.array([[float(np.ravel(e)[0]) for e in row] for row in x])
    mod.np = _NP()
    return mod


def make_cadence():
    """cadence.Fisher (unmodified reference function) on analytic spectroscopic inputs; 6 shells, every integer l in [60, 1500]."""
    from . import halofit as ohf
    hf = refload.load('halofit')
    prep = refload.load('prep_Llls')
    cad = load_cadence()
    k, p, eng, cosmo, P = _pk_setup()
    Dg = lambda z: float(eng.Dgrow(z=z))
    ref_pk = ohf.ModiPk(k, p, Dg, cosmo, P['sig_8'], halofit_fn=hf.halofit).interps()
    _, Lm, _nm = prep.prep_Llls(60, 60, 1500, log10=False)
    T = cadence_inputs()
    out = {'k': k, 'plin': p, 'L': Lm, 'zs': T['zs'], 'tNs': T['tNs'], 'tNp': T['tNp'], 'fsky': T['fsky'], 'fover': T['fover']}
    cad.dz = 0.1
    t = time.time()
    for name, sources in (('lenses', False), ('sources', True)):
        res = cad.Fisher(ref_pk, Lm, T['zs'], T['tNs'], T['tNp'], T['pz'], T['bz'], fsky=T['fsky'], fover=T['fover'], sources=sources, printit=False)
        out['out_' + name] = np.asarray(res, dtype=np.float64)
    out['lenses_bz'] = np.array([float(cad.lenses_bz(z)) for z in T['zs']])
    out['Cij'] = cad.Cij(ref_pk, Lm[:50].astype(np.float64), 0.45, 0.55)
    print('cadence.Fisher reference: %.1f s' % (time.time() - t), out['out_lenses'][:2])
    np.savez_compressed(os.path.join(GOLD, 'cadence.npz'), **out)


def cadence_inputs():
    """Analytic stand-ins for get_desiCz() / desi_eff (cadence.py:25-41, schmittfull_bz.py): shared with the GPU test."""
    g = lambda z0, s: (lambda z: np.exp(-(np.asarray(z, dtype=np.float64) - z0) ** 2 / (2. * s * s)) / (np.sqrt(2. * np.pi) * s))
    return {'zs': np.arange(0.3, 0.85, 0.1), 'tNs': 1500., 'tNp': 1.2e5, 'pz': g(0.6, 0.3), 'bz': (lambda z: 1.2 + 0.6 * z),
            'fsky': 5060. / 41253., 'fover': 0.05}


def variance_inputs():
    """Analytic inputs of the variance / ellipse goldens, shared with tests/test_variances_gpu.py."""
    g = lambda z0, s: (lambda z: np.exp(-(np.asarray(z, dtype=np.float64) - z0) ** 2 / (2. * s * s)) / (np.sqrt(2. * np.pi) * s))
    return {'pz': g(2.96, 0.24), 'bz': (lambda z: 1. + z), 'bz2': (lambda z: 0.8 + 0.5 * z), 'nbar': 2181.0, 'fsky': 0.4,
            'zmin': 0.96, 'zmax': 4.96, 'rho': (lambda L: 0.8 * np.exp(-np.asarray(L, dtype=np.float64) / 3000.)),
            'nkk': (lambda L: 2.0e-8 * (1. + (np.asarray(L, dtype=np.float64) / 1500.) ** 2.)),
            'covs': [np.array([[2.0e-4, 3.0e-4], [3.0e-4, 9.0e-3]]), np.array([[4.0, -1.5], [-1.5, 1.0]]),
                     np.array([[1.0e-2, 0.0], [0.0, 1.0e-2]])]}


def make_variances():
    """The variance wrappers (Cgg.py:70-83, Ckg.py:60-87, cmb/lensing.py:54-66), prep_camb.Clxy (:128-137), pmh.Pmm/Pmh/Phh
    (:26-47), Nkk.prep_lmaxmask (cmb/Nkk.py:13-26) and fisher_contour.plot_ellipse (:15-111), all run UNMODIFIED."""
    from . import halofit as ohf
    from lbgcmb_b200.synthetic import cmb_cl_tables
    hf = refload.load('halofit')
    Cgg, Ckg, lensing, Nkk = refload.load('Cgg'), refload.load('Ckg'), refload.load('lensing'), refload.load('Nkk')
    prep_camb, pmh, fcont = refload.load('prep_camb'), refload.load('pmh'), refload.load('fisher_contour')
    prep = refload.load('prep_Llls')
    k, p, eng, cosmo, P = _pk_setup()
    Dg = lambda z: float(eng.Dgrow(z=z))
    ref_pk = ohf.ModiPk(k, p, Dg, cosmo, P['sig_8'], halofit_fn=hf.halofit).interps()
    T = variance_inputs()
    _, Llls, nmodes = prep.prep_Llls(60, 50, 5000, log10=True)
    Ld = Llls.astype(np.float64)
    zmin, zmax, fsky, nbar = T['zmin'], T['zmax'], T['fsky'], T['nbar']
    nkk = T['nkk'](Ld)
    out = {'k': k, 'plin': p, 'Llls': Llls, 'nkk': nkk}
    t = time.time()
    cgg_full = Cgg.Cgg(ref_pk, Ld, zmin, zmax, T['pz'], T['bz'], zeff=False)
    out['cgg_full'] = cgg_full
    out['var_cgg'] = Cgg.var_Cgg(Ld, zmin, zmax, T['pz'], T['bz'], nbar, fsky, Pk_interps=ref_pk)
    out['var_cgg_sv'] = Cgg.var_Cgg(Ld, zmin, zmax, T['pz'], T['bz'], nbar, fsky, Pk_interps=ref_pk, samplevar_lim=True)
    out['var_cgg_given'] = Cgg.var_Cgg(Ld, zmin, zmax, T['pz'], T['bz'], nbar, fsky, cgg=cgg_full)
    li, ni = None, None
    vk = lambda **kw: Ckg.var_Ckg(ref_pk, li, ni, Ld, zmin, zmax, T['pz'], T['bz'], nbar, fsky, nkk=nkk, **kw)
    out['var_ckg'] = vk()
    out['var_ckg_full'] = vk(zeff=False)
    out['var_ckg_rho'] = vk(rho=T['rho'])
    out['var_ckg_msl'] = vk(modesample_lim=True)
    out['var_ckg_ngg'] = vk(ngg=3.0e-9 * np.ones_like(Ld))
    out['var_ckk'] = lensing.var_Ckk(Ld, fsky, nkk, Pk_interps=ref_pk)
    out['var_ckk_sv'] = lensing.var_Ckk(Ld, fsky, nkk, Pk_interps=ref_pk, samplevar_lim=True)
    print('variance wrappers: %.1f s' % (time.time() - t))
    # nkk=None: the wrapper runs Nkk itself (two multipoles; thetab / DeltaT / iterative forwarded)
    t = time.time()
    ell, lens, nolens = cmb_cl_tables(5300)
    li, ni = on0.cl_interps(ell, lens), on0.cl_interps(ell, nolens)
    L2 = np.array([100., 2000.])
    out['L2'] = L2
    out['var_ckg_nkknone'] = Ckg.var_Ckg(ref_pk, li, ni, L2, zmin, zmax, T['pz'], T['bz'], nbar, fsky, thetab=1.4, DeltaT=7.0,
                                         iterative=True, pickleit=False)
    print('var_Ckg(nkk=None): %.1f s' % (time.time() - t))
    # Clxy on integer and fractional multipoles
    x = np.array([2., 10.5, 100., 1000.25, 2999.5, 5000.])
    out['clxy_x'] = x
    for a in ('TT', 'EE', 'BB', 'TE'):
        out['clxy_' + a] = prep_camb.Clxy(li, x.copy(), a, 1.4, 7.0)
    # P(k,z) helpers
    kk = np.array([1.0e-3, 0.02, 0.1, 0.5, 2.0, 30.0])
    out['pk_k'], out['pk_z'] = kk, 1.3
    out['pmm_lin'] = pmh.Pmm(ref_pk, kk, 1.3, type='linear')
    out['pmm_nlin'] = pmh.Pmm(ref_pk, kk, 1.3)
    out['pmh'] = pmh.Pmh(ref_pk, T['bz'], kk, 1.3)
    out['phh'] = pmh.Phh(ref_pk, T['bz'], kk, 1.3)
    out['phh2'] = pmh.Phh(ref_pk, T['bz'], kk, 1.3, bz2=T['bz2'])
    # prep_lmaxmask on values either side of every cut
    nu = np.array([1., 2., 2999.5, 3000., 3000.5, 4999., 5000., 5000.5, 10., 10.])
    nv = np.array([10., 10., 10., 10., 10., 10., 10., 10., 1.5, 5001.])
    m = Nkk.prep_lmaxmask(nu, nv)
    out['mask_nu'], out['mask_nv'] = nu, nv
    out['mask_keys'] = np.array(sorted(m))
    out['masks'] = np.array([m[key] for key in sorted(m)])
    # Fisher ellipses
    for i, cov in enumerate(T['covs']):
        for tag, ml in (('68', 0.68), ('95', 0.95), ('none', None)):
            out['ell_%d_%s' % (i, tag)] = fcont.plot_ellipse(cov=cov.copy(), mass_level=ml, data_out=True, theta_num=1000,
                                                              x_cent=0.8, y_cent=3.9)
    out['ell_axes'] = fcont.plot_ellipse(semimaj=2.0, semimin=0.5, phi=0.3, theta_num=500, data_out=True)
    np.savez_compressed(os.path.join(GOLD, 'variances.npz'), **out)
    print('variances golden written:', sorted(out))


EH_CASES = [(0.292, 0.0468, 0.0, 1, 0.708, 0.69, 0.0), (0.3106, 0.04898, 0.0, 3, 0.6894, 0.677, 3.0),
            (0.31, 0.049, 0.0045, 3, 0.69, 0.68, 1.0), (0.25, 0.04, 0.02, 1, 0.70, 0.72, 0.5)]


def make_eh():
    """T_cb(k), T_cbnu(k) from the compiled reference (oracle/_ref/libeh_ref.so <- EisensteinHu/EisensteinAndHu.c, float32)."""
    from . import eh
    if not eh.ref_available():
        raise RuntimeError('run `make -C oracle` first')
    k = np.exp(np.linspace(np.log(1e-4), np.log(50.), 400))
    out = {'k': k, 'cases': np.array(EH_CASES)}
    for i, args in enumerate(EH_CASES):
        cb, cbnu = eh.ref_transfer(args, k)
        out['tf_cb_%d' % i], out['tf_cbnu_%d' % i] = cb, cbnu
    np.savez_compressed(os.path.join(GOLD, 'eh.npz'), **out)
    print('eh golden:', [float(out['tf_cb_%d' % i][200]) for i in range(len(EH_CASES))])


MAKERS = {'n0': make_n0, 'halofit': make_halofit, 'limber': make_limber, 'fisher': make_fisher, 'modi': make_modi,
          'modi_lensing': make_modi_lensing, 'cadence': make_cadence, 'eh': make_eh, 'variances': make_variances}

if __name__ == '__main__':
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or sorted(MAKERS)
    for w in which:
        MAKERS[w]()
