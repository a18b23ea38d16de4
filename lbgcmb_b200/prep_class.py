"""CLASS C_l files as inputs of the N0 kernels; mirrors reference prep_class.py:4-50 (`prep_classCls`) and gives the same files
to the root-flavour N0 (cmb/Nkk.py) as integer-l tables.

CLASS writes l, TT, EE, TE, BB, phiphi, TPhi, Ephi with the l(l+1)/2pi prefactor and l = 2..lmax (prep_class.py:12-25,
modi/nkk.py:168-192).  The reference hard-codes dat/RunPB/RunPB02_cl{,_lensed}.dat, which its tree does not ship; here the two
paths are arguments with the reference's as defaults.  The splines are table-backed `modi.nkk.interpolate` objects (FITPACK
k = 3 interpolating spline on unit-spaced knots, extrapolating: ext = 0 as prep_class.py:17), evaluated on the GPU."""
import numpy as np

from lbgcmb_b200.DetectorNoise import DetectorNoise

COLS = ['ell', 'TT', 'EE', 'TE', 'BB', 'PP', 'TP', 'EP']
SPECTRA = {'TT': 0, 'EE': 1, 'TE': 2, 'BB': 3}


def read_class(path):
    """{'ell', 'TT', ..., 'EP', 'kk'}: raw C_l (prefactor l(l+1)/2pi removed, prep_class.py:23-24) and
    C_l^kk = (l(l+1))^2 C_l^phiphi / 4 (:26)."""
    raw = np.loadtxt(path, ndmin=2)
    if raw.shape[1] < len(COLS):
        raise ValueError('%s: expected the %d CLASS columns %s, found %d' % (path, len(COLS), ' '.join(COLS), raw.shape[1]))
    data = {name: raw[:, i].copy() for i, name in enumerate(COLS)}
    ell = data['ell']
    for x in COLS[1:]:
        data[x] /= ell * (ell + 1.) / (2. * np.pi)
    data['kk'] = 0.25 * (ell * (ell + 1.)) ** 2. * data['PP']
    return data


def prep_classCls(lensed="dat/RunPB/RunPB02_cl_lensed.dat", unlensed="dat/RunPB/RunPB02_cl.dat"):
    """(lensCl_interps, nolensCl_interps) keyed TT, EE, TE, BB.  As in the reference the LENSED splines carry the detector
    noise of a thetab = 1', DeltaT = 1 uK-arcmin experiment (prep_class.py:37) and the unlensed ones do not (:46)."""
    from lbgcmb_b200.modi.nkk import interpolate
    data = read_class(lensed)
    lens = {x: interpolate(data['ell'], data[x] + DetectorNoise(data['ell'], 1., 1., x)) for x in SPECTRA}
    data = read_class(unlensed)
    nolens = {x: interpolate(data['ell'], data[x]) for x in SPECTRA}
    return lens, nolens


def class_cl_tables(lensed="dat/RunPB/RunPB02_cl_lensed.dat", unlensed="dat/RunPB/RunPB02_cl.dat"):
    """(ell, lensed[4][n], unlensed[4][n]) on l = 0..lmax in the order TT, EE, BB, TE with C_0 = C_1 = 0: the table layout of
    CAMB.get_Cls (prep_camb.py:84-89) that `prep_camb.cl_interps_from_tables` and Nkk.Nkk consume, built from CLASS files
    (no detector noise: Nkk adds its own).  The two files are cut to their common lmax (CLASS writes 500 more unlensed l)."""
    out = []
    lmax = None
    for path in (lensed, unlensed):
        d = read_class(path)
        ell = d['ell']
        if ell[0] < 0 or np.any(np.abs(ell - np.round(ell)) > 0) or np.any(np.diff(ell) != 1.0):
            raise ValueError('%s: multipoles must be consecutive integers' % path)
        tab = np.zeros((4, int(ell[-1]) + 1))
        for i, x in enumerate(('TT', 'EE', 'BB', 'TE')):
            tab[i, int(ell[0]):] = d[x]
        out.append(tab)
        lmax = tab.shape[1] if lmax is None else min(lmax, tab.shape[1])
    return np.arange(lmax, dtype=np.float64), out[0][:, :lmax].copy(), out[1][:, :lmax].copy()
