"""Gaussian p(z); mirrors reference Gaussian_pz.py:5-29 (host-side).  The normalising integral is scipy-1.2.1
`simps(ps, dx=dz)` on 1000 (even) samples -> quadrature.simps_weights."""
import numpy as np

from lbgcmb_b200.quadrature import simps_weights


def Gaussian(z, z0, sigma):
    arg = -(z - z0) ** 2.
    arg /= 2. * sigma ** 2.
    return np.exp(arg)


def Gaussian_pz(z, z0=2.96, sigma=0.24):
    dz = 0.01
    zs = np.arange(0., 10., dz)
    ps = Gaussian(zs, z0, sigma)
    norm = np.dot(simps_weights(zs.size, dz), ps)
    return Gaussian(z, z0, sigma) / norm
