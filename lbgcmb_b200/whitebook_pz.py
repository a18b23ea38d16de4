"""LSST redshift distributions and number densities; mirrors reference whitebook_pz.py:4-98 (host-side input providers:
`const_pz`, `whitebook_pz`, `whitebook_dpdz`, `Ng`, `lsst_chang_pz`).  A few closed-form expressions per call, as in the
reference; the forecast kernels consume their samples on the z grids."""
import numpy as np


def const_pz(z, zlo=0.0, zhi=1.0):
    if isinstance(z, float):
        if (zlo < z) & (z < zhi):
            return 1. / (zhi - zlo)
        return 0.0
    z = np.asarray(z)
    result = np.ones_like(z) / (zhi - zlo)
    result[z < zlo] = 0.0
    result[z > zhi] = 0.0
    return result


def whitebook_pz(z, ilim=25.3):
    z0 = 0.0417 * ilim - 0.744                      # LSST science book (3.8)
    zz = z / z0
    return 0.5 * zz ** 2. * np.exp(-zz) / z0


def whitebook_dpdz(z, ilim):
    z0 = 0.0417 * ilim - 0.744
    zz = z / z0
    return 0.5 * (2. * zz - zz ** 2.) * np.exp(-zz) / (z0 ** 2.)


def Ng(ilim, fmask=0.12):
    """Galaxies per sq. deg. (arXiv:1809.01669); Y1: 24.1, Y10: 25.3."""
    result = 42.9 * (1. - fmask) * 10. ** (0.359 * (ilim - 25.0))
    result *= 60. * 60.
    return result


_CHANG = {(False, 24.1): (0.26, 0.94), (False, 25.3): (0.28, 0.90), (True, 24.1): (0.13, 0.78), (True, 25.3): (0.11, 0.68)}


def lsst_chang_pz(z, ilim=24.1, source=False):
    dz = 0.01
    zs = np.arange(0., 10., dz)
    if source is False and (False, ilim) not in _CHANG:
        raise ValueError("\n\nRequested ilim not available for LSST (Chang) p(z). \n\n")
    if (bool(source), ilim) not in _CHANG:
        # the reference leaves lengthscale / exponent unbound here and dies with UnboundLocalError (whitebook_pz.py:72-87)
        raise UnboundLocalError("local variable 'lengthscale' referenced before assignment")
    lengthscale, exponent = _CHANG[(bool(source), ilim)]
    norm = zs * zs * np.exp(-(zs / lengthscale) ** exponent)
    result = z * z * np.exp(-(z / lengthscale) ** exponent)
    result /= np.sum(norm * dz)
    return result
