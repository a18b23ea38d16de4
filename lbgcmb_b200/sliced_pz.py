"""p(z) renormalised on [zmin, zmax]; mirrors reference sliced_pz.py:4-27 (host-side)."""
import numpy as np


def sliced_pz(zz, zmin, zmax, survey_pz):
    dz = 0.001
    zs = np.arange(zmin, zmax, dz)
    ps = survey_pz(zs)
    norm = np.trapezoid(ps, zs)
    result = survey_pz(zz) / norm
    result[zz < zmin] = 0.0
    result[zz > zmax] = 0.0
    # constant linear spacing of the z array (to six decimal places)
    assert len(set(np.around(np.ediff1d(zz), 6))) == 1
    dz = list(set(np.around(np.ediff1d(zz), 6)))[0]
    result /= (np.sum(result) * dz)
    return result
