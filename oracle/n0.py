"""CPU restatement of the Hu-Okamoto N0 lensing-reconstruction noise (TEST INFRASTRUCTURE).

root flavour : cmb/Nkk.py:31-184 (Nkk), cmb/Nkk.py:13-26 (prep_lmaxmask),
               prep_camb.py:128-137 (Clxy), cmb/DetectorNoise.py:8-32 (DetectorNoise)
modi flavour : modi/nkk.py:35-92 (N_fisher), :94-146 (N_fisherTE), :194-261 (do_noise / do_noiseTE)

Pinned against the unmodified reference by oracle/make_golden.py -> tests/golden/n0_*.npz
(tests/test_oracle_pinning.py).  The root flavour is evaluated one L at a time (the reference
builds the whole (phi, l1, L) cube, ~0.5 GB per L); every per-node operation and the order of the
two Simpson passes is the reference's, so results are bit-identical per L.
"""
import numpy as np
from scipy.interpolate import interp1d, InterpolatedUnivariateSpline

from .simps121 import simps

SPECTRA = ('TT', 'EE', 'BB', 'TE')  # prep_camb.py:84


def detector_noise(ell, thetab, DeltaT, kind, Tcmb):
    """cmb/DetectorNoise.py:8-32; thetab [arcmin], DeltaT [uK-arcmin]."""
    DeltaT = DeltaT * 10. ** -6.
    arcmin2radians = np.pi / (60. * 180.)
    DeltaT = DeltaT * arcmin2radians
    thetab = thetab * arcmin2radians
    DeltaP = np.sqrt(2.) * DeltaT
    if kind == 'TT':
        return (DeltaT / Tcmb) ** 2. * np.exp(ell * (ell + 1.) * thetab ** 2. / (8. * np.log(2.)))
    if kind in ('EE', 'BB', 'EB'):
        return (DeltaP / Tcmb) ** 2. * np.exp(ell * (ell + 1.) * thetab ** 2. / (8. * np.log(2.)))
    if kind == 'TE':
        return 0.0
    raise ValueError("Noise estimate is not available for %s combination." % kind)


def cl_interps(ell, table4):
    """prep_camb.py:88-89: cubic interp1d per spectrum, 0 outside the table."""
    return {x: interp1d(ell, table4[i], kind='cubic', bounds_error=False, fill_value=0.0)
            for i, x in enumerate(SPECTRA)}


def root_grids(lmax=5000., Nphi=200, Nell=10000):
    """cmb/Nkk.py:54-62 -> (ell[9999], pphi[201], dell, dphi)."""
    dphi = 2. * np.pi / Nphi
    dell = lmax / Nell
    ell = np.arange(1.0, lmax + dell, dell)
    pphi = np.arange(0.0, 2. * np.pi + dphi, dphi)
    return ell, pphi, dell, dphi


def nkk_root(lens, nolens, Ls, terms=('TT', 'TE', 'EE', 'EB'), thetab=1., DeltaT=1., iterative=False,
             Tcmb=2.786, lmax=5000., Nphi=200, Nell=10000, cos_phi=None, sin_phi=None):
    """N0_kk(L) as cmb/Nkk.py:54-172.  `lens` / `nolens`: dict name -> callable (as cl_interps()).

    cos_phi / sin_phi default to np.cos / np.sin of the reference phi grid; they can be supplied so a
    golden vector made on one machine replays bit-identically on another (libm differences in the last
    bit move the |u| > 3000 mask at l1 == 3000, cmb/Nkk.py:17).
    """
    ell, pphi, dell, dphi = root_grids(lmax, Nphi, Nell)
    if cos_phi is None:
        cos_phi, sin_phi = np.cos(pphi), np.sin(pphi)
    l = np.broadcast_to(ell[None, :], (pphi.size, ell.size))
    c = cos_phi[:, None]
    s = sin_phi[:, None]

    def tot(name, x):  # Clxy, prep_camb.py:128-137
        return lens[name](x) + detector_noise(x, thetab, DeltaT, name, Tcmb)

    out = np.zeros(len(Ls))
    for iL, L in enumerate(Ls):
        L = float(L)
        ux = l * c
        uy = l * s
        vx = L - ux
        vy = -uy
        phi_u = np.arctan2(uy, ux)
        phi_v = np.arctan2(vy, vx)
        phi_u[phi_u < 0.] += 2. * np.pi
        phi_v[phi_v < 0.] += 2. * np.pi
        phi_uv = phi_u - phi_v
        nu = (ux ** 2. + uy ** 2.) ** 0.5
        nv = (vx ** 2. + vy ** 2.) ** 0.5
        Lx = ux + vx
        dLu = ux * Lx
        dLv = vx * Lx
        m3 = np.ones_like(nu)
        m5 = np.ones_like(nu)
        m3[nu > 3000] = 0.0
        m3[nv > 3000] = 0.0
        m5[nu > 5000] = 0.0
        m5[nv > 5000] = 0.0
        m3[nv < 2] = 0.0
        m5[nv < 2] = 0.0
        result = np.zeros_like(ux)
        if 'TT' in terms:
            fa = nolens['TT'](nu) * dLu + nolens['TT'](nv) * dLv
            Fa = fa / (2. * tot('TT', nu) * tot('TT', nv))
            result += fa * Fa * m3
        if 'TE' in terms:
            c2 = np.cos(2. * phi_uv)
            TT1, TT2 = tot('TT', nu), tot('TT', nv)
            EE1, EE2 = tot('EE', nu), tot('EE', nv)
            TE1, TE2 = tot('TE', nu), tot('TE', nv)
            fa_uv = nolens['TE'](nu) * c2 * dLu + nolens['TE'](nv) * dLv
            fa_vu = nolens['TE'](nu) * dLu + nolens['TE'](nv) * c2 * dLv
            num = EE1 * TT2 * fa_uv - TE1 * TE2 * fa_vu
            denom = TT1 * EE2 * EE1 * TT2 - (TE1 * TE2) ** 2.
            result += fa_uv * (num / denom) * m3
        if 'EE' in terms:
            fa = (nolens['EE'](nu) * dLu + nolens['EE'](nv) * dLv) * np.cos(2. * phi_uv)
            Fa = fa / (2. * tot('EE', nu) * tot('EE', nv))
            result += fa * Fa * m5
        if 'EB' in terms:
            fa = (nolens['EE'](nu) * dLu - nolens['BB'](nv) * dLv) * np.sin(2. * phi_uv)
            Fa = fa / (tot('EE', nu) * tot('BB', nv))
            if iterative:
                result += fa * Fa * m5 * 2.5
            else:
                result += fa * Fa * m5
        if 'BB' in terms:
            fa = (nolens['BB'](nu) * dLu + nolens['BB'](nv) * dLv) * np.cos(2. * phi_uv)
            Fa = fa / (2. * tot('BB', nu) * tot('BB', nv))
            result += fa * Fa * m5
        if 'TB' in terms:
            raise KeyError('TB')  # reference indexes nolensCl_interps['TB'], which get_Cls never builds
        result *= l
        r = simps(result, dx=dell, axis=0)   # over phi, labelled dell (cmb/Nkk.py:166)
        r = simps(r, dx=dphi, axis=0)        # over l1, labelled dphi
        r /= (2. * np.pi) ** 2.
        out[iL] = (1. / r) * (L * (L + 1.) / 2.) ** 2.
    return out


# ---------------------------------------------------------------- modi flavour
def modi_N_fisher(L, l, icl1, icl2, icln1, icln2, mode, nphi=200):
    """modi/nkk.py:35-92."""
    phis = np.linspace(0, 2 * np.pi, nphi)
    cosphi = np.cos(phis)
    cl1 = icl1(l)
    cln1 = icln1(l)
    Llcosphi = L * l.reshape(-1, 1) * cosphi.reshape(1, -1)
    l2 = (L ** 2 + (l ** 2).reshape(-1, 1) - 2 * Llcosphi) ** 0.5
    cl2 = icl2(l2)
    cln2 = icln2(l2)
    num = (Llcosphi * cl1.reshape(-1, 1) + (L ** 2 - Llcosphi) * cl2) ** 2.
    den = cln1.reshape(-1, 1) * cln2
    if mode == "TT":
        den *= 2.
    else:
        sinphi12 = L * np.sin(phis).reshape(1, -1) / l2
        cosphi12 = (L * cosphi.reshape(1, -1) - l.reshape(-1, 1)) / l2
        if mode == "EB":
            num *= (2 * sinphi12 * cosphi12) ** 2.
        elif mode == "EE":
            num *= (cosphi12 ** 2. - sinphi12 ** 2.) ** 2.
            den *= 2.
    frac = num / den
    frac[l2 > l.max()] = 0.
    frac[l2 < l.min()] = 0.
    int1 = l * simps(frac, phis, axis=-1)
    return simps(int1, l)


def modi_N_fisherTE(L, l, iclte, iclnt, iclne, iclnte, nphi=200):
    """modi/nkk.py:94-146."""
    phis = np.linspace(0, 2 * np.pi, nphi)
    cosphi = np.cos(phis)
    Llcosphi = L * l.reshape(-1, 1) * cosphi.reshape(1, -1)
    l2 = (L ** 2. + (l ** 2.).reshape(-1, 1) - 2. * Llcosphi) ** 0.5
    Ll2cosphi = L ** 2. - Llcosphi
    sinphi12 = L * np.sin(phis).reshape(1, -1) / l2
    cosphi12 = (L * cosphi.reshape(1, -1) - l.reshape(-1, 1)) / l2
    cos2phi12 = cosphi12 ** 2. - sinphi12 ** 2.
    cl1te, cl2te = iclte(l).reshape(-1, 1), iclte(l2)
    cl1ntt, cl2ntt = iclnt(l).reshape(-1, 1), iclnt(l2)
    cl1nee, cl2nee = iclne(l).reshape(-1, 1), iclne(l2)
    cl1nte, cl2nte = iclnte(l).reshape(-1, 1), iclnte(l2)
    fl1l2 = Llcosphi * cl1te * cos2phi12 + Ll2cosphi * cl2te
    fl2l1 = Llcosphi * cl1te + Ll2cosphi * cl2te * cos2phi12
    num = cl1nee * cl2ntt * fl1l2 - cl1nte * cl2nte * fl2l1
    den = cl1ntt * cl2nee * cl1nee * cl2ntt - (cl1nte * cl2nte) ** 2.
    frac = fl1l2 * (num / den)
    frac[l2 > l.max()] = 0.
    frac[l2 < l.min()] = 0.
    int1 = l * simps(frac, phis, axis=-1)
    return simps(int1, l)


def modi_do_noise(ell_u, cl_u, ell_l, cl_l, mode, thetab, delT, lat, Tcmb=2.786, n_l=10000):
    """modi/nkk.py:194-261 on in-memory tables.  cl_u / cl_l: dict name -> raw C_l on ell_u / ell_l.

    mode in {'TT','EE','EB','TE'}.  Returns N0(lat) for that estimator alone.
    """
    lat = np.asarray(lat, dtype=np.float64)
    IUS = InterpolatedUnivariateSpline
    if mode == 'TE':
        lint = np.linspace(1., 3000., n_l)
        icl3 = IUS(ell_u, cl_u['TE'])
        icln1 = IUS(ell_l, cl_l['TT'] + detector_noise(ell_l, thetab, delT, 'TT', Tcmb))
        icln2 = IUS(ell_l, cl_l['EE'] + detector_noise(ell_l, thetab, delT, 'EE', Tcmb))
        icln3 = IUS(ell_l, cl_l['TE'])
        noise = np.array([modi_N_fisherTE(L, lint, icl3, icln1, icln2, icln3) for L in lat])
    else:
        a1, a2 = mode[0] * 2, mode[1] * 2
        lint = np.linspace(1., 5000. if mode in ('EB', 'EE') else 3000., n_l)
        icl1, icl2 = IUS(ell_u, cl_u[a1]), IUS(ell_u, cl_u[a2])
        icln1 = IUS(ell_l, cl_l[a1] + detector_noise(ell_l, thetab, delT, a1, Tcmb))
        icln2 = IUS(ell_l, cl_l[a2] + detector_noise(ell_l, thetab, delT, a2, Tcmb))
        noise = np.array([modi_N_fisher(L, lint, icl1, icl2, icln1, icln2, mode) for L in lat])
    noise = (2 * np.pi) ** 2. / noise
    return noise * (lat * (lat + 1)) ** 2. / 4.
