"""p(z) of a dropout sample from the reference's table; mirrors dropouts/schechter/get_pz.py:8-14."""
import os

import numpy as np
from scipy.interpolate import interp1d


def get_pz(type='g'):
    root = os.environ['LBGCMB']
    file = root + '/dropouts/schechter/dat/pz/pz_%s.dat' % type
    x, y = np.loadtxt(file, unpack=True)
    return interp1d(x, y, kind='linear', copy=True, bounds_error=False, fill_value=0.0, assume_sorted=False)
