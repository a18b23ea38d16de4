"""Quadrature weight vectors.  The CUDA kernels take weights, never "a rule"; these builders reproduce the rules the
reference calls: scipy 1.2.1 `simps(even='avg')` (pinned in condaenv_minimal.yml:42; Cgg.py:68, Ckg.py:57,
cmb/lensing.py:43, modi/nkk.py:88-90, modi/halofit.py:42) and `np.trapz` (modi/lensing.py:137,167,179)."""
import numpy as np


def _add_panels(w, h, start, stop, scale):
    """3-point (unequal spacing) Simpson panels on samples start, start+2, ... < stop, scaled."""
    i = np.arange(start, stop, 2)
    if i.size == 0:
        return
    h0, h1 = h[i], h[i + 1]
    hsum = h0 + h1
    r = h0 / h1
    np.add.at(w, i, scale * hsum / 6.0 * (2 - 1.0 / r))
    np.add.at(w, i + 1, scale * hsum / 6.0 * hsum * hsum / (h0 * h1))
    np.add.at(w, i + 2, scale * hsum / 6.0 * (2 - r))


def simps_weights_x(x):
    """w such that sum(w*y) == scipy-1.2.1 simps(y, x) (even='avg' for an even number of samples)."""
    x = np.asarray(x, dtype=np.float64)
    n = x.size
    h = np.diff(x)
    w = np.zeros(n)
    if n % 2 == 1:
        _add_panels(w, h, 0, n - 2, 1.0)
    else:
        _add_panels(w, h, 0, n - 3, 0.5)      # Simpson on the first n-1 samples + trapezoid on the last interval
        _add_panels(w, h, 1, n - 2, 0.5)      # trapezoid on the first interval + Simpson on the last n-1 samples
        w[-1] += 0.25 * h[-1]
        w[-2] += 0.25 * h[-1]
        w[0] += 0.25 * h[0]
        w[1] += 0.25 * h[0]
    return w


def simps_weights(n, dx=1.0):
    """w such that sum(w*y) == scipy-1.2.1 simps(y, dx=dx)."""
    w = np.zeros(n)

    def panels(start, stop, scale):
        i = np.arange(start, stop, 2)
        np.add.at(w, i, scale * dx / 3.0)
        np.add.at(w, i + 1, scale * 4.0 * dx / 3.0)
        np.add.at(w, i + 2, scale * dx / 3.0)

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


def trapz_weights_x(x):
    """w such that sum(w*y) == np.trapz(y, x)."""
    x = np.asarray(x, dtype=np.float64)
    h = np.diff(x)
    w = np.zeros(x.size)
    w[:-1] += 0.5 * h
    w[1:] += 0.5 * h
    return w


def romberg(f, a, b, tol=1.48e-8, rtol=1.48e-8, divmax=10):
    """scipy-1.2.1 `romberg(f, a, b, vec_func=True)` for a vectorised integrand: the rule the reference's growth factor is
    built with (modi/cosmology.py:107; removed from scipy >= 1.15).  Host-side helper, <= 1025 ordinates per call."""
    span = b - a
    acc = 0.5 * (f(a) + f(b))
    prev = [span * acc]
    out = prev[0]
    panels = 1
    for level in range(1, divmax + 1):
        panels *= 2
        half = panels / 2
        h = float(b - a) / half
        acc = acc + np.sum(f((a + 0.5 * h) + h * np.arange(half)), axis=0)
        cur = [span * acc / panels]
        for k in range(level):
            p4 = 4.0 ** (k + 1)
            cur.append((p4 * cur[k] - prev[k]) / (p4 - 1.0))
        out = cur[level]
        err = abs(out - prev[level - 1])
        if err < tol or err < rtol * abs(out):
            break
        prev = cur
    return out
