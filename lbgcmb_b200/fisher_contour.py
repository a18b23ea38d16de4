"""Fisher ellipse of a 2x2 covariance; same call surface as reference fisher_contour.py:15 (`plot_ellipse`).

Host only (one 2x2 SVD, a 20001-point chi^2_2 table lookup, `theta_num` samples).  Importing this module needs neither
matplotlib nor latexify; a plot is attempted only when the caller does not ask for `data_out`."""
import numpy as np
from scipy.stats import chi2


def _contour_scale(mass_level):
    """sqrt of the chi^2 (2 dof) distance enclosing `mass_level`; table lookup as fisher_contour.py:76-81."""
    if mass_level is None:
        return np.sqrt(2.279)
    grid = np.linspace(0, 20, 20001)
    miss = np.abs(chi2.cdf(grid, df=2) - mass_level)
    return np.sqrt(grid[np.flatnonzero(miss == miss.min())[0]])


def ellipse_from_cov(cov, mass_level=0.68):
    """(semi-major, semi-minor, position angle) of the `mass_level` contour; fisher_contour.py:65-89."""
    vecs, vals, _ = np.linalg.svd(cov)
    lead = vecs[0]
    if lead[0] < 0:
        lead *= -1
    k = _contour_scale(mass_level)
    a, b = np.sqrt(vals[0]) * k, np.sqrt(vals[1]) * k
    ang = np.arccos(np.dot(lead, np.array([1, 0])))
    if lead[1] < 0 and ang > 0:
        ang *= -1
    return a, b, ang


def ellipse_points(a, b, ang, x0=0, y0=0, n=1e3):
    """2 x n samples of the rotated ellipse; fisher_contour.py:92-108 (including its unit-radius normalisation)."""
    t = np.linspace(0, 2 * np.pi, int(n))
    ct, st = np.cos(t), np.sin(t)
    rad = 1 / np.sqrt(ct ** 2 + st ** 2)
    unit = np.array([rad * ct, rad * st])
    rot = np.array([[np.cos(ang), -np.sin(ang)], [np.sin(ang), np.cos(ang)]])
    pts = np.dot(np.dot(rot, np.array([[a, 0], [0, b]])), unit)
    pts[0] += x0
    pts[1] += y0
    return pts


def plot_ellipse(semimaj=1, semimin=1, phi=0, x_cent=0, y_cent=0, theta_num=1e3, ax=None,
                 plot_kwargs=None, fill=False, fill_kwargs=None, data_out=False, cov=None, mass_level=0.68):
    if cov is not None:
        semimaj, semimin, phi = ellipse_from_cov(cov, mass_level)
    data = ellipse_points(semimaj, semimin, phi, x_cent, y_cent, theta_num)
    if data_out == True:  # noqa: E712 (the reference tests identity with True)
        return data
    import pylab as pl     # only reached when a plot is requested
    made = ax is None
    if made:
        fig, ax = pl.subplots()
    if plot_kwargs is None:
        ax.plot(data[0], data[1], color='b', linestyle='-', lw=2)
    else:
        ax.plot(data[0], data[1], **plot_kwargs)
    if fill == True:  # noqa: E712
        ax.fill(data[0], data[1], **fill_kwargs)
    if made:
        return fig
