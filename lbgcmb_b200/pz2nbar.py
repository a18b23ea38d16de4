"""Integrated p(z) and number-density unit conversion; mirrors reference pz2nbar.py:6-31 (host-side)."""
import numpy as np


def pz_slice(zmin, zmax, pz):
    dz = 0.001
    zz = np.arange(zmin, zmax, dz)
    return np.trapezoid(pz(zz), zz, dx=dz)


def nbar_convert(nbar, unit='arcmin'):
    if unit == 'arcmin':
        return nbar / 60. / 60.
    elif unit == 'str':
        deg_str = np.pi / 180.
        return nbar / deg_str ** 2.
    else:
        raise ValueError("Requested unit is not available for dropout p(z).")
