"""scipy 1.2.1 `scipy.integrate.romberg` restated (TEST INFRASTRUCTURE).

The reference pins scipy 1.2.1 (condaenv_minimal.yml:42) and calls `romberg(kernel, a, b, vec_func=True)` with the
default tolerances for the growth factor (modi/cosmology.py:107); scipy >= 1.15 removed the function.  Algorithm of
scipy/integrate/quadrature.py (1.2.1): trapezoid sums on 1, 2, 4, ... 2^divmax panels, Richardson extrapolation of each
new row, stop when the last two diagonal entries agree to tol (absolute) or rtol (relative)."""
import numpy as np


def _difftrap(function, interval, numtraps):
    """Sum of the *new* ordinates when the panel count doubles to `numtraps` (half-weighted ends for one panel)."""
    if numtraps == 1:
        return 0.5 * (function(interval[0]) + function(interval[1]))
    numtosum = numtraps / 2
    h = float(interval[1] - interval[0]) / numtosum
    lox = interval[0] + 0.5 * h
    points = lox + h * np.arange(numtosum)
    return np.sum(function(points), axis=0)


def _romberg_diff(b, c, k):
    tmp = 4.0 ** k
    return (tmp * c - b) / (tmp - 1.0)


def romberg(function, a, b, args=(), tol=1.48e-8, rtol=1.48e-8, show=False, divmax=10, vec_func=False):
    if np.isinf(a) or np.isinf(b):
        raise ValueError("Romberg integration only available for finite limits.")
    if vec_func:
        vfunc = lambda x: function(x, *args)
    else:
        def vfunc(x):
            if np.isscalar(x):
                return function(x, *args)
            return np.array([function(xi, *args) for xi in np.asarray(x)])
    n = 1
    interval = [a, b]
    intrange = b - a
    ordsum = _difftrap(vfunc, interval, n)
    result = intrange * ordsum
    last_row = [result]
    for i in range(1, divmax + 1):
        n *= 2
        ordsum += _difftrap(vfunc, interval, n)
        row = [intrange * ordsum / n]
        for k in range(i):
            row.append(_romberg_diff(last_row[k], row[k], k + 1))
        result = row[i]
        lastresult = last_row[i - 1]
        err = abs(result - lastresult)
        if err < tol or err < rtol * abs(result):
            break
        last_row = row
    return result
