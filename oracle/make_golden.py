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


MAKERS = {'n0': make_n0, 'halofit': make_halofit, 'limber': make_limber}

if __name__ == '__main__':
    os.makedirs(GOLD, exist_ok=True)
    which = sys.argv[1:] or sorted(MAKERS)
    for w in which:
        MAKERS[w]()
