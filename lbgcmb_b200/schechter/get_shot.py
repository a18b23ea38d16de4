"""Cumulative surface density of a dropout sample at a magnitude limit; mirrors dropouts/schechter/get_shot.py:6-23."""
import os

import numpy as np


def get_shot(type='g', mlim=25.5):
    files = {'BX': 'BXDrop.dat', 'u': 'MalkanDrop.dat', 'g': 'gDrop.dat', 'r': 'rDrop.dat', 'qso': 'qso.dat'}
    file = files[type]
    root = os.environ['LBGCMB']
    fpath = root + '/dropouts/schechter/dat/' + file
    ms, ns = np.loadtxt(fpath, unpack=True)                  # per sq. deg.
    if mlim == None:                                         # noqa: E711 (reference spelling)
        return ms, ns
    index = (np.abs(ms - mlim) == np.abs(ms - mlim).min())
    return ns[index][0]                                      # per sq. deg.
