"""scipy 1.2.1 ``scipy.integrate.simps`` restated (TEST INFRASTRUCTURE, see oracle/__init__.py).

The reference pins scipy 1.2.1 (condaenv_minimal.yml:42) and calls ``simps`` at
Cgg.py:68, Ckg.py:57, cmb/lensing.py:43, cmb/Nkk.py:166-167, modi/nkk.py:88-90,142-144,
modi/halofit.py:42.  For an even number of samples that version averages
"Simpson on the first N-1 points + trapezoid on the last interval" with
"trapezoid on the first interval + Simpson on the last N-1 points" (even='avg');
scipy >= 1.14 replaced that with a different end correction, so it cannot serve as oracle.

Two forms are provided:
  simps(y, x=None, dx=1, axis=-1)   array-in / array-out, same operation order as scipy 1.2.1
  simps_weights(n, dx) / simps_weights_x(x)   the equivalent explicit weight vectors the CUDA kernels
                                      consume ("kernels take weights, never a rule").
"""
import numpy as np


def _composite(y, start, stop, x, dx, axis):
    """Sum of 3-point Simpson panels starting at samples start, start+2, ... < stop."""
    y = np.moveaxis(y, axis, -1)
    y0 = y[..., start:stop:2]
    y1 = y[..., start + 1:stop + 1:2]
    y2 = y[..., start + 2:stop + 2:2]
    if x is None:
        return np.sum(dx / 3.0 * (y0 + 4 * y1 + y2), axis=-1)
    h = np.diff(np.moveaxis(x, axis, -1), axis=-1)
    h0 = h[..., start:stop:2]
    h1 = h[..., start + 1:stop + 1:2]
    hsum = h0 + h1
    hprod = h0 * h1
    h0divh1 = h0 / h1
    tmp = hsum / 6.0 * (y0 * (2 - 1.0 / h0divh1) + y1 * hsum * hsum / hprod + y2 * (2 - h0divh1))
    return np.sum(tmp, axis=-1)


def simps(y, x=None, dx=1, axis=-1, even='avg'):
    if even != 'avg':
        raise ValueError("only even='avg' (the scipy 1.2.1 default the reference relies on) is restated")
    y = np.asarray(y)
    nd = y.ndim
    n = y.shape[axis]
    if x is not None:
        x = np.asarray(x)
        if x.ndim == 1:
            shp = [1] * nd
            shp[axis] = x.shape[0]
            x = x.reshape(shp)
        elif x.ndim != nd:
            raise ValueError("If given, shape of x must be 1-d or the same as y.")
        if x.shape[axis] != n:
            raise ValueError("If given, length of x along axis must be the same as y.")
    if n % 2 == 1:
        return _composite(y, 0, n - 2, x, dx, axis)
    ym = np.moveaxis(y, axis, -1)
    first_dx = last_dx = dx
    if x is not None:
        xm = np.moveaxis(x, axis, -1)
        last_dx = xm[..., -1] - xm[..., -2]
        first_dx = xm[..., 1] - xm[..., 0]
    val = 0.5 * last_dx * (ym[..., -1] + ym[..., -2])
    result = _composite(y, 0, n - 3, x, dx, axis)
    val = val + 0.5 * first_dx * (ym[..., 1] + ym[..., 0])
    result = result + _composite(y, 1, n - 2, x, dx, axis)
    return result / 2.0 + val / 2.0


def simps_weights(n, dx=1.0):
    """w with sum(w*y) == simps(y, dx=dx) up to summation order, for n uniform samples."""
    return _uniform_weights(n, dx)


def _uniform_weights(n, dx):
    w = np.zeros(n)

    def panels(start, stop, scale):
        for i in range(start, stop, 2):
            w[i] += scale * dx / 3.0
            w[i + 1] += scale * 4.0 * dx / 3.0
            w[i + 2] += scale * dx / 3.0

    if n % 2 == 1:
        panels(0, n - 2, 1.0)
    else:
        panels(0, n - 3, 0.5)
        panels(1, n - 2, 0.5)
        w[-1] += 0.25 * dx
        w[-2] += 0.25 * dx
        w[0] += 0.25 * dx
        w[1] += 0.25 * dx
    return w


def simps_weights_x(x):
    """w with sum(w*y) == simps(y, x) (unequal-spacing three-point formula of scipy 1.2.1)."""
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    h = np.diff(x)
    w = np.zeros(n)

    def panels(start, stop, scale):
        for i in range(start, stop, 2):
            h0, h1 = h[i], h[i + 1]
            hsum = h0 + h1
            r = h0 / h1
            w[i] += scale * hsum / 6.0 * (2 - 1.0 / r)
            w[i + 1] += scale * hsum / 6.0 * hsum * hsum / (h0 * h1)
            w[i + 2] += scale * hsum / 6.0 * (2 - r)

    if n % 2 == 1:
        panels(0, n - 2, 1.0)
    else:
        panels(0, n - 3, 0.5)
        panels(1, n - 2, 0.5)
        w[-1] += 0.25 * h[-1]
        w[-2] += 0.25 * h[-1]
        w[0] += 0.25 * h[0]
        w[1] += 0.25 * h[0]
    return w


def trapz_weights_x(x):
    """w with sum(w*y) == np.trapz(y, x)."""
    x = np.asarray(x, dtype=np.float64)
    h = np.diff(x)
    w = np.zeros(x.size)
    w[:-1] += 0.5 * h
    w[1:] += 0.5 * h
    return w
