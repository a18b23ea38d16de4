import numpy as np
from scipy.special import hyp2f1

_C_KMS = 299792.458


class _Q:
    """Quantity-like: only .value is used by the reference."""

    def __init__(self, v):
        self.value = v

    # growth_rate.py:17 forms (H(0) / H(z)) ** 2 * a ** -3 on Quantities before taking .value
    def __truediv__(self, o):
        return _Q(self.value / (o.value if isinstance(o, _Q) else o))

    def __mul__(self, o):
        return _Q(self.value * (o.value if isinstance(o, _Q) else o))

    __rmul__ = __mul__

    def __pow__(self, e):
        return _Q(self.value ** e)


class FlatLambdaCDM:
    """Flat LCDM with Tcmb0 = 0 (matter + Lambda only), as astropy builds it for cosmo.py:6."""

    def __init__(self, H0, Om0, Ob0=None, name=None):
        self._H0 = float(H0)
        self.Om0 = float(Om0)
        self.Ob0 = Ob0
        self.Ode0 = 1.0 - self.Om0
        self.Onu0 = 0.0
        self.name = name

    def efunc(self, z):
        z = np.asarray(z, dtype=np.float64)
        return np.sqrt(self.Om0 * (1.0 + z) ** 3 + self.Ode0)

    def H(self, z):
        return _Q(self._H0 * self.efunc(z))

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
        # astropy 3.1 closed form for flat LCDM without radiation (hypergeometric), [Mpc]
        z = np.asarray(z, dtype=np.float64)
        s = ((1.0 - self.Om0) / self.Om0) ** (1.0 / 3)
        pref = (_C_KMS / self._H0) / np.sqrt(s * self.Om0)
        return _Q(pref * (self._T(s) - self._T(s / (1.0 + z))))
