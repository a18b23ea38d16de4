"""Background cosmology; mirrors reference cosmo.py:6 (`cosmo = FlatLambdaCDM(H0, Om0, Ob0)`).

astropy is not a dependency: `FlatLambdaCDM` below provides exactly the members the hot path touches
(H(z).value, comoving_distance(z).value, efunc, Om, Ode, w, Onu0, Om0) with astropy-3.1's closed form for the flat,
radiation-free case (Tcmb0 = 0): D_C = (c/H0) [T(s) - T(s/(1+z))] / sqrt(s Om0), T(x) = 2 sqrt(x) 2F1(1/6,1/2;7/6;-x^3)
(SURVEY.md N2).  Host-side: <= 2000 doubles per call, as in the reference."""
import numpy as np
from scipy.special import hyp2f1

from lbgcmb_b200.params import get_params

C_KMS = 299792.458


class _Quantity(object):
    def __init__(self, value):
        self.value = value


class FlatLambdaCDM(object):
    def __init__(self, H0, Om0, Ob0=None, name=None):
        self.H0 = _Quantity(float(H0))
        self.Om0 = float(Om0)
        self.Ob0 = Ob0
        self.Ode0 = 1.0 - self.Om0
        self.Onu0 = 0.0
        self.name = name

    def efunc(self, z):
        z = np.asarray(z, dtype=np.float64)
        return np.sqrt(self.Om0 * (1.0 + z) ** 3 + self.Ode0)

    def H(self, z):
        return _Quantity(self.H0.value * self.efunc(z))

    def Om(self, z):
        z = np.asarray(z, dtype=np.float64)
        return self.Om0 * (1.0 + z) ** 3 / self.efunc(z) ** 2

    def Ode(self, z):
        return self.Ode0 / self.efunc(z) ** 2

    def w(self, z):
        return -1.0

    @staticmethod
    def _T(x):
        return 2.0 * np.sqrt(x) * hyp2f1(1.0 / 6, 1.0 / 2, 7.0 / 6, -x ** 3)

    def comoving_distance(self, z):
        z = np.asarray(z, dtype=np.float64)
        s = ((1.0 - self.Om0) / self.Om0) ** (1.0 / 3)
        pref = (C_KMS / self.H0.value) / np.sqrt(s * self.Om0)
        return _Quantity(pref * (self._T(s) - self._T(s / (1.0 + z))))


params = get_params()
cosmo = FlatLambdaCDM(H0=100. * params['h_100'], Om0=params['om_m'], Ob0=params['om_b'])
