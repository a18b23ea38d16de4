"""CPU restatement of HALOFIT and the modi P(k,z) provider (TEST INFRASTRUCTURE).

modi/halofit.py:7-100 (_get_spec, main sigma_8 branch), :102-214 (halofit); modi/cosmology.py:158-182 (pkalin, pkanlin).
Pinned against the unmodified reference `halofit` by oracle/make_golden.py -> tests/golden/halofit.npz.
"""
import numpy as np
from scipy.interpolate import InterpolatedUnivariateSpline as _spline

from .simps121 import simps as _simps


def get_spec(k, delta_k, sigma_8):
    """modi/halofit.py:35-98, branch 0.6 < sigma_8 < 1 (the only one the forecast cosmologies reach)."""
    if not (sigma_8 < 1.0 and sigma_8 > 0.6):
        raise NotImplementedError('oracle restates the 0.6 < sigma_8 < 1 branch only')
    lnr = np.linspace(np.log(0.1), np.log(10.0), 500)
    lnsig = np.empty(500)
    lnk = np.log(k)
    for i, r in enumerate(lnr):
        R = np.exp(r)
        integrand = delta_k * np.exp(-(k * R) ** 2)
        lnsig[i] = np.log(_simps(integrand, lnk))
    r_of_sig = _spline(lnsig[::-1], lnr[::-1], k=5)
    rknl = 1.0 / np.exp(r_of_sig(0.0))
    sig_of_r = _spline(lnr, lnsig, k=5)
    dev1, dev2 = sig_of_r.derivatives(np.log(1.0 / rknl))[1:3]
    return rknl, -dev1 - 3.0, -dev2


def halofit(k, delta_k, sigma_8, z, cosmo, takahashi=True):
    """modi/halofit.py:102-214."""
    rknl, neff, rncur = get_spec(k, delta_k, sigma_8)
    mask = k > 0.005
    plin = delta_k[mask]
    k = k[mask]
    omegamz = cosmo.Om(z)
    omegavz = cosmo.Ode(z)
    w = cosmo.w(z)
    fnu = cosmo.Onu0 / cosmo.Om0
    if takahashi:
        a = 10 ** (1.5222 + 2.8553 * neff + 2.3706 * neff ** 2 + 0.9903 * neff ** 3 + 0.2250 * neff ** 4 +
                   - 0.6038 * rncur + 0.1749 * omegavz * (1 + w))
        b = 10 ** (-0.5642 + 0.5864 * neff + 0.5716 * neff ** 2 + - 1.5474 * rncur + 0.2279 * omegavz * (1 + w))
        c = 10 ** (0.3698 + 2.0404 * neff + 0.8161 * neff ** 2 + 0.5869 * rncur)
        gam = 0.1971 - 0.0843 * neff + 0.8460 * rncur
        alpha = np.abs(6.0835 + 1.3373 * neff - 0.1959 * neff ** 2 + - 5.5274 * rncur)
        beta = (2.0379 - 0.7354 * neff + 0.3157 * neff ** 2 + 1.2490 * neff ** 3 + 0.3980 * neff ** 4 - 0.1682 * rncur +
                fnu * (1.081 + 0.395 * neff ** 2))
        xmu = 0.0
        xnu = 10 ** (5.2105 + 3.6902 * neff)
    else:
        a = 10 ** (1.4861 + 1.8369 * neff + 1.6762 * neff ** 2 + 0.7940 * neff ** 3 + 0.1670 * neff ** 4 + - 0.6206 * rncur)
        b = 10 ** (0.9463 + 0.9466 * neff + 0.3084 * neff ** 2 + - 0.94 * rncur)
        c = 10 ** (-0.2807 + 0.6669 * neff + 0.3214 * neff ** 2 - 0.0793 * rncur)
        gam = 0.8649 + 0.2989 * neff + 0.1631 * rncur
        alpha = np.abs(1.3884 + 0.3700 * neff - 0.1452 * neff ** 2)
        beta = (0.8291 + 0.9854 * neff + 0.3401 * neff ** 2)
        xmu = 10 ** (-3.5442 + 0.1908 * neff)
        xnu = 10 ** (0.9589 + 1.2857 * neff)
    if np.abs(1 - omegamz) > 0.01:
        f1a, f2a, f3a = omegamz ** -0.0732, omegamz ** -0.1423, omegamz ** 0.0725
        f1b, f2b, f3b = omegamz ** -0.0307, omegamz ** -0.0585, omegamz ** 0.0743
        frac = omegavz / (1 - omegamz)
        if takahashi:
            f1, f2, f3 = f1b, f2b, f3b
        else:
            f1 = frac * f1b + (1 - frac) * f1a
            f2 = frac * f2b + (1 - frac) * f2a
            f3 = frac * f3b + (1 - frac) * f3a
    else:
        f1 = f2 = f3 = 1.0
    y = k / rknl
    ph = a * y ** (f1 * 3) / (1 + b * y ** f2 + (f3 * c * y) ** (3 - gam))
    ph = ph / (1 + xmu / y + xnu * y ** -2) * (1 + fnu * (0.977 - 18.015 * (cosmo.Om0 - 0.3)))
    plinaa = plin * (1 + fnu * 47.48 * k ** 2 / (1 + 1.5 * k ** 2))
    pq = plin * (1 + plinaa) ** beta / (1 + plinaa * alpha) * np.exp(-y / 4.0 - y ** 2 / 8.0)
    out = delta_k.copy()
    out[mask] = pq + ph
    return out


class ModiPk(object):
    """modi/cosmology.py:158-182 on in-memory tables.  `Dgrow(z)` is an input (host-side growth, same callable the CUDA
    path uses); `cosmo` supplies Om(z), Ode(z), w(z), Onu0, Om0.  `halofit_fn` lets the pinning test swap in the
    unmodified reference `halofit`."""

    def __init__(self, klin, plin, Dgrow, cosmo, sig8=0.819, halofit_fn=None):
        self.klin, self.plin, self.Dgrow, self.cosmo, self.sig8 = klin, plin, Dgrow, cosmo, sig8
        self.halofit_fn = halofit_fn or halofit
        self._cache = {}

    def pkalin(self, z):
        a = 1. / (1. + z)
        return self.klin, self.plin * self.Dgrow(z) ** 2.

    def pkanlin(self, z):
        key = float(z)
        if key not in self._cache:
            a = 1. / (1. + z)
            if z > 4:
                self._cache[key] = self.pkalin(z)[1]
            else:
                pdiml = self.plin * self.klin ** 3 / (2. * np.pi ** 2)
                pgive = pdiml * self.Dgrow(z) ** 2.
                pndim = self.halofit_fn(k=self.klin, delta_k=pgive, sigma_8=self.sig8, z=1. / a - 1., cosmo=self.cosmo, takahashi=True)
                self._cache[key] = pndim * (2. * np.pi ** 2) / self.klin ** 3
        return self.klin, self._cache[key]

    def interps(self):
        """{'camb_lin','camb_nlin'} -> callable (z, k[]) as pmh.py:18-24 expects; np.interp in k, zero outside the table."""
        return {'camb_lin': lambda z, k: np.interp(k, self.klin, self.pkalin(z)[1], left=0.0, right=0.0),
                'camb_nlin': lambda z, k: np.interp(k, self.klin, self.pkanlin(z)[1], left=0.0, right=0.0)}
