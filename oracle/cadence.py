"""CPU restatement of the LSST-lens clustering-redshift forecast (TEST INFRASTRUCTURE).

lenses_bz, Fisher      cadence/cadence.py:43-47, :64-191 (as written: global `dz`, LCUT of the last shell, running `Interim`)
lsst_chang_pz          whitebook_pz.py:53-93          growth_factor   growth_rate.py:9-30
Pinned against the unmodified reference function (loaded by oracle/make_golden.py:load_cadence) -> tests/golden/cadence.npz.
"""
from collections import OrderedDict

import numpy as np

from . import limber as ol

KMAX = 0.3


def lsst_chang_pz(z, ilim=25.3, source=False):
    zs = np.arange(0., 10., 0.01)
    lengthscale, exponent = {(False, 24.1): (0.26, 0.94), (False, 25.3): (0.28, 0.90),
                             (True, 24.1): (0.13, 0.78), (True, 25.3): (0.11, 0.68)}[(bool(source), ilim)]
    norm = zs * zs * np.exp(-(zs / lengthscale) ** exponent)
    return z * z * np.exp(-(z / lengthscale) ** exponent) / np.sum(norm * 0.01)


def growth_factor(bg, aa):
    def rate(a):
        z = 1. / a - 1.
        return (bg.p['om_m'] * (bg.H(0.) / bg.H(z)) ** 2. * a ** -3.) ** 0.545
    avals = np.arange(0.001, 1., 0.001)
    integrand = (rate(avals) - 1.) / avals
    norm = np.exp(np.trapezoid(integrand, avals))
    return aa * np.exp(np.trapezoid(integrand[avals <= aa], avals[avals <= aa])) / norm


def lenses_bz(bg, zee):
    return 0.95 / growth_factor(bg, 1. / (1. + zee))


def Fisher(bg, Pk, Llls, zs, tNs, tNp, pz, bz, fsky=0.1, fover=0.0, sources=False, dz=0.1):
    result = OrderedDict()
    to_str = lambda n: n / (np.pi / 180.) ** 2.
    lbz = lambda z: lenses_bz(bg, z)
    for zee in zs:
        LMAX = np.ceil(KMAX * bg.chi(zee) - 0.5).astype('int')
        LCUT = np.where(Llls <= LMAX)
        bs = bz(zee)
        Ns = to_str(tNs) * pz(zee) * dz
        bp = 1. + zee if sources else lbz(zee)
        Np = to_str(tNp) * lsst_chang_pz(zee, 25.3, False) * dz
        wp, ws = Np, Ns
        wps = fover * np.min([wp, ws])
        cgg = ol.Cgg(bg, Pk, Llls, zee - dz / 2., zee + dz / 2., pz, bz,
                     survey_pz2=lambda z: lsst_chang_pz(z, 25.3, sources), bz2=lbz, zeff=True)
        r = {'Ns': Ns, 'Np': Np, 'num': (bp * bs * cgg) ** 2., 'A00z': cgg * (bp * Np) ** 2. + wp,
             'A0i': bp * Np * bs * Ns * cgg + wps, 'Aii': (bs * Ns) ** 2. * cgg + ws, 'A0i_j': bp * bs * Ns * cgg}
        result[zee] = r
    A00 = sum(r['A00z'] for r in result.values())
    sshift = sum(r['A0i'] ** 2. / r['Aii'] for r in result.values())
    S = A00 / (A00 - sshift)
    for r in result.values():
        r['Fii'] = r['Ns'] * fsky * np.sum((2. * Llls[LCUT] + 1.) * (r['num'] / A00)[LCUT])
        r['ferr_ii'] = np.sqrt(1. / r['Fii']) / r['Np']
        r['r'] = r['A0i'] / np.sqrt(A00 * r['Aii'])
    F = np.zeros((len(zs), len(zs)))
    for i, ri in enumerate(result.values()):
        Interim = (2. * Llls + 1) * S * ri['A0i_j'] * ri['A0i_j'] / ri['Aii'] / A00
        F[i, i] = fsky * np.sum(Interim[LCUT])
        for j, rj in enumerate(result.values()):
            Interim = Interim + (2. * Llls + 1) * fsky * 2. * S ** 2. * ri['r'] * rj['r'] \
                * np.sqrt(1. / ri['Aii'] / rj['Aii']) * ri['A0i_j'] * rj['A0i_j'] / A00
            F[i, j] += np.sum(Interim[LCUT])
    diFish = np.diag(np.linalg.inv(F))
    return np.array([[z, r['Ns'], r['Np'], 100. * r['ferr_ii'], 100. * np.sqrt(diFish[i]) / r['Np']]
                     for i, (z, r) in enumerate(result.items())])
