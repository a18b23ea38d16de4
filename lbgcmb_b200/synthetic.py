"""Deterministic synthetic inputs for tests and bench.py.

CAMB is absent (and unpinned in the reference), so CMB spectra and the linear P(k) are analytic
stand-ins with realistic magnitudes; they are *inputs*, identical for the CUDA path and the oracle.
"""
import numpy as np

# RunPB cosmology, reference params.py:6
PARAMS = {'om_m': 0.292, 'om_b': 0.0468, 'om_L': 0.708, 'h_100': 0.69, 'sig_8': 0.819,
          'Tcmb': 2.786, 'zscatter': 1100, 'ns': 0.9688, 'As': 2.187e-9}


def cmb_cl_tables(n_ell=10151):
    """Raw (no l(l+1)/2pi), dimensionless C_l on integer l = 0..n_ell-1, as prep_camb.py:70-100 holds them.

    Returns (ell, lensed[4, n_ell], unlensed[4, n_ell]) in the order TT, EE, BB, TE.
    """
    ell = np.arange(n_ell, dtype=np.float64)
    with np.errstate(divide='ignore', invalid='ignore'):
        ll1 = ell * (ell + 1.0)
        tt = 1.0e-9 / ll1 * np.exp(-(ell / 1400.0) ** 1.2) * (1.0 + 0.3 * np.cos(ell / 95.0))
        ee = 3.0e-11 / ll1 * np.exp(-(ell / 1800.0) ** 1.2) * (1.0 + 0.4 * np.cos(ell / 95.0 + 1.3))
        bb = 2.0e-13 / ll1 * np.exp(-(ell / 2500.0) ** 1.1)
        te = 1.0e-10 / ll1 * np.exp(-(ell / 1600.0) ** 1.2) * np.cos(ell / 150.0)
    lensed = np.stack([tt, ee, bb, te])
    lensed[:, :2] = 0.0
    unlensed = 1.02 * lensed
    unlensed[2] = 0.0  # no primordial B modes (prep_camb.py:84-85 comment)
    return ell, lensed, unlensed


def eh_like_plin(n_k=2000, kmin=1.5e-5, kmax=100.0, params=PARAMS):
    """Linear P(k) at z=0 on a log-uniform k grid [h/Mpc] -> (k, P [(Mpc/h)^3]), sigma8-normalised.

    BBKS-shaped transfer function; layout follows dropouts/Martin/dat/thy/Planck18_pklin.txt.
    """
    k = np.exp(np.linspace(np.log(kmin), np.log(kmax), n_k))
    gamma = params['om_m'] * params['h_100'] * np.exp(-params['om_b'] * (1.0 + np.sqrt(2 * params['h_100']) / params['om_m']))
    q = k / gamma
    t = np.log(1.0 + 2.34 * q) / (2.34 * q) * (1 + 3.89 * q + (16.1 * q) ** 2 + (5.46 * q) ** 3 + (6.71 * q) ** 4) ** -0.25
    p = k ** params['ns'] * t ** 2
    x = 8.0 * k
    w = 3.0 * (np.sin(x) - x * np.cos(x)) / x ** 3
    w = np.where(x < 1e-3, 1.0 - x * x / 10.0, w)
    lnk = np.log(k)
    integrand = p * k ** 3 / (2 * np.pi ** 2) * w * w
    s2 = np.sum(0.5 * (integrand[1:] + integrand[:-1]) * np.diff(lnk))
    p *= params['sig_8'] ** 2 / s2
    return k, p


def gaussian_pz(z0=2.96, sigma=0.24):
    """Un-normalised Gaussian p(z) callable; shape of Gaussian_pz.py:5-9 (u-dropouts default)."""
    def pz(z):
        z = np.asarray(z, dtype=np.float64)
        return np.exp(-(z - z0) ** 2 / (2.0 * sigma ** 2)) / (np.sqrt(2 * np.pi) * sigma)
    return pz


def linz_bz(z):
    """b(z) = 1 + z, pmh.py:12."""
    return 1.0 + np.asarray(z, dtype=np.float64)


DROPOUTS = {  # (z0, sigma, nbar [deg^-2]) -- SURVEY.md 8(d)
    'u': (2.96, 0.24, 2181.0), 'g': (3.8, 0.3, 5409.0), 'r': (4.9, 0.3, 1332.0), 'i': (5.9, 0.3, 300.0)}


def write_class_files(clfile, clsfile, n_unlensed=5799, n_lensed=5299):
    """CLASS-format text files as modi/nkk.py:168-192 reads them: columns l, TT, EE, TE, BB, phiphi, TPhi, Ephi for
    l = 2.., spectra multiplied by l(l+1)/2pi; the unlensed file runs 500 multipoles further (modi/nkk.py:186)."""
    ell, lens, nolens = cmb_cl_tables(n_unlensed + 2)
    for path, tab, n in ((clfile, nolens, n_unlensed), (clsfile, lens, n_lensed)):
        l = ell[2:n + 2]
        f = l * (l + 1.) / (2. * np.pi)
        pp = 1.0e-7 / (l * (l + 1.)) ** 2 * np.exp(-l / 800.)
        cols = [l, tab[0, 2:n + 2] * f, tab[1, 2:n + 2] * f, tab[3, 2:n + 2] * f, tab[2, 2:n + 2] * f, pp * f, 0. * l, 0. * l]
        np.savetxt(path, np.array(cols).T, fmt='%.17e')
