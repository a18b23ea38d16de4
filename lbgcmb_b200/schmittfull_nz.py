"""Tracer samples of Schmittfull & Seljak (2017); mirrors reference schmittfull_nz.py:12-64 (`ss_pz`, `load_samples`,
`get_ss17_samples`).  The reference reads `dat/...` relative to the working directory (it is run from $LBGCMB); here the same
files are looked up under $LBGCMB when they are not found relative to the cwd.  Host-side file readers, as in the reference."""
import os

import numpy as np
import pandas as pd
from scipy.interpolate import interp1d

from lbgcmb_b200.quadrature import simps_weights
from lbgcmb_b200.schmittfull_bz import get_allbz
from lbgcmb_b200.whitebook_pz import const_pz


def _dat(name):
    if os.path.isfile(os.path.join('dat', name)):
        return os.path.join('dat', name)
    return os.path.join(os.environ['LBGCMB'], 'dat', name)


def ss_pz():
    """LSST dN/dz of SS17 -> (p(z) interp1d, galaxies per sq. deg.); schmittfull_nz.py:12-28 (9e6-node Simpson, scipy-1.2.1 rule)."""
    data = np.loadtxt(_dat('schmittfull_nbar.dat'))
    nz_interp = interp1d(data[:, 0], data[:, 1], kind='nearest', bounds_error=False, fill_value=0.0, assume_sorted=False)
    zs = np.arange(0., 9., 1.e-6)
    dndz = nz_interp(zs)
    norm = float(np.dot(simps_weights(zs.size, zs[1] - zs[0]), dndz))          # [deg^-2]
    pz = interp1d(zs, dndz / norm, kind='nearest', bounds_error=False, fill_value=0.0, assume_sorted=False)
    return pz, 3600. * norm


def load_samples():
    return pd.read_csv(_dat('ss17_tracers.dat'), names=['zmin', 'zmax', 'N', 'Surveys'], sep=r'\s+', comment='#')


def get_ss17_samples(nolsst=True, printit=False):
    dframe = load_samples()
    surveys = list(dframe['Surveys'].values)
    zlos = dframe['zmin'].values.astype('float')
    zhis = dframe['zmax'].values.astype('float')
    all_bs = get_allbz()
    ns = dframe['N'].values.astype('float')
    bs = [all_bs[x] for x in surveys]
    ps = []
    for x in zip(zlos, zhis):
        # as in the reference (:45-47) the lambda closes over the loop VARIABLE: after the loop every p(z) is the top-hat of
        # the LAST row of dat/ss17_tracers.dat (0.6 < z < 1.9).  Kept, because the rho_L the reference ships were made this way.
        pz = lambda z: const_pz(z, zlo=x[0], zhi=x[1])
        ps += [np.vectorize(pz)]
    if not nolsst:
        pz, norm = ss_pz()
        ns += [norm]                               # on the numpy array this ADDS norm to every entry (reference :53), it does not append
        ps += [pz]
        bs += [lambda z: (1. + z)]
        surveys += ['LSST']
    if printit:
        print('\n\nAvailable surveys: ' + ''.join('  %s;' % s for s in surveys) + '\n\n')
    return ns, ps, bs, surveys
