"""Linear growth from f = Omega_m(a)^0.545; mirrors reference growth_rate.py:9-30 (host-side, used by pmh.invG_bz)."""
import numpy as np

from lbgcmb_b200.cosmo import cosmo
from lbgcmb_b200.params import get_params

params = get_params()


def growth_rate(a):
    gamma = 0.545
    z = 1. / a - 1.
    om_ma = params['om_m'] * (cosmo.H(0).value / cosmo.H(z).value) ** 2. * a ** -3.
    return om_ma ** gamma


def _growth_factor(aa):
    avals = np.arange(0.001, 1., 0.001)
    integrand = (growth_rate(avals) - 1.) / avals
    norm = np.exp(np.trapezoid(integrand, avals))
    result = aa * np.exp(np.trapezoid(integrand[avals <= aa], avals[avals <= aa]))
    return result / norm


growth_factor = np.vectorize(_growth_factor)
