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


_MISSING = -98.9          # entries at or below -99 mark (z, m) cells without a measurement


def _bias_table(path):
    """(row labels, column labels, body) of a labelled table file: first row / column hold the labels, the corner is unused;
    cells marked missing come back masked."""
    raw = np.loadtxt(path)
    body = np.where(raw[1:, 1:] < _MISSING, np.nan, raw[1:, 1:])
    return raw[1:, 0], raw[0, 1:], ma.masked_invalid(body)


class drop_bz(object):
    """Measured dropout bias b(z, m_lim) with asymmetric errors from the table pair <type>.dat / <type>_err.dat
    (dropouts/bz/py/get_bz.py:34-111 reads the same files): attributes zs, ms (the tabulated redshifts and magnitude limits), bs
    [n_z, n_m] (masked where not measured), el / eu (lower / upper errors: the two column halves of the error file) and the table
    limits zlo, zhi, mlo, mhi."""

    def __init__(self, type='harikane4'):
        stem = os.path.join(os.environ['LBGCMB'], 'dropouts', 'bz', 'dat', type)
        self.zs, self.ms, self.bs = _bias_table(stem + '.dat')
        _, _, err = _bias_table(stem + '_err.dat')
        n_m = self.ms.size
        self.el, self.eu = err[:, :n_m], err[:, n_m:]
        self.zlo, self.zhi = self.zs.min(), self.zs.max()
        self.mlo, self.mhi = self.ms.min(), self.ms.max()

    def _nearest(self, nodes, x):
        d = np.abs(nodes - x)
        return d == d.min()                  # boolean selector: ties select both neighbours, as the reference's lookup does

    def get_bzm(self, z, m, printit=False):
        """(b, lower error, upper error) at the tabulated node nearest to (z, m); UserWarning outside the table."""
        if not (self.zlo <= z <= self.zhi and self.mlo <= m <= self.mhi):
            raise UserWarning('\n\nHarikane bias not available at z=%.2lf' % z)
        iz, im = self._nearest(self.zs, z), self._nearest(self.ms, m)
        out = self.bs[iz, im], self.el[iz, im], self.eu[iz, im]
        if printit:
            print(*out)
        return out
