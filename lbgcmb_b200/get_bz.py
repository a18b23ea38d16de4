"""Linear bias of dropout samples; mirrors dropouts/bz/py/get_bz.py:19-32 (`bz_fitmodel`, `bz_callmodel`) and the table
container :34-111 (`drop_bz`: Harikane et al. bias tables under $LBGCMB/dropouts/bz/dat).  Host-side, as in the reference."""
import os

import numpy as np
import numpy.ma as ma

from lbgcmb_b200.growth_rate import growth_factor


def bz_fitmodel(z, A, B):
    aa = 1. / (1. + z)
    DD = growth_factor(aa)                                   # noqa: F841 (evaluated and unused in the reference too, :21)
    return A / aa + B / aa ** 2.


@np.vectorize
def bz_callmodel(z, mlim):
    mm = mlim - 25.
    A = -0.978 * mm + 0.1110
    B = 0.1158 * mm + 0.1692
    return bz_fitmodel(z, A, B)


class drop_bz():
    def __init__(self, type='harikane4'):
        root = os.environ['LBGCMB'] + '/dropouts/bz/dat/'
        har = np.loadtxt(root + '%s.dat' % type)
        hare = np.loadtxt(root + '%s_err.dat' % type)
        zs = har[:, 0][1:]
        mags = har[0, :][1:]
        har[har < -98.9] = np.nan
        hare[hare < -98.9] = np.nan
        har = ma.masked_invalid(har)[1:, 1:]
        hare = ma.masked_invalid(hare)[1:, 1:]
        cut = len(mags)
        self.zs, self.ms, self.bs = zs, mags, har
        self.el, self.eu = hare[:, :cut], hare[:, cut:]
        self.zlo, self.zhi = self.zs.min(), self.zs.max()
        self.mlo, self.mhi = self.ms.min(), self.ms.max()

    def get_bzm(self, z, m, printit=False):
        if (z >= self.zlo) & (z <= self.zhi) & (m >= self.mlo) & (m <= self.mhi):
            zindex = np.abs(self.zs - z) == np.abs(self.zs - z).min()
            mindex = np.abs(self.ms - m) == np.abs(self.ms - m).min()
            if printit:
                print(self.bs[zindex, mindex], self.el[zindex, mindex], self.eu[zindex, mindex])
            return self.bs[zindex, mindex], self.el[zindex, mindex], self.eu[zindex, mindex]
        raise UserWarning('\n\nHarikane bias not available at z=%.2lf' % z)
