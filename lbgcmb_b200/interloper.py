"""(sigma8, b1) Fisher forecast and interloper bias; mirrors reference interloper.py:23-74 (`get_allCls`, `get_ClsCov`,
`dp_DL`) and restates its script body :200-260 as `fisher_bias` (same arithmetic, same order of the sums).

The reference's `get_allCls` reads lensCl_interps, nolensCl_interps, thetab, DeltaT and iterative from module globals that
only exist when the file runs as a script (interloper.py:25-38 vs :84-110).  Here the same names are module
attributes (`configure(...)`) and may also be given as keyword arguments.
Device work: N0, the Limber spectra, the per-L Gaussian covariance (lbg_gaussian_cov), the per-L 3x3 solves and the
Fisher / bias sums over L (lbg_fisher_accum).  The final 2x2 inverse is numpy, as in the reference (:256)."""
import numpy as np
from numpy.linalg import inv

from lbgcmb_b200 import fisher as _fisher
from lbgcmb_b200.Cgg import Cgg, var_Cgg, Ngg
from lbgcmb_b200.Ckg import Ckg, var_Ckg
from lbgcmb_b200.Nkk import Nkk
from lbgcmb_b200.lensing import Ckk, var_Ckk
from lbgcmb_b200.params import get_params

params = get_params()

# the script-level globals of interloper.py:84-110
lensCl_interps = None
nolensCl_interps = None
thetab = 1.
DeltaT = 1.
iterative = False


def configure(**kw):
    """Set the module globals `get_allCls` reads (lensCl_interps, nolensCl_interps, thetab, DeltaT, iterative)."""
    g = globals()
    for k, v in kw.items():
        if k not in ('lensCl_interps', 'nolensCl_interps', 'thetab', 'DeltaT', 'iterative'):
            raise TypeError('unknown setting %s' % k)
        g[k] = v


def get_allCls(Pk_interps, Llls, nbar, fsky, zmin, zmax, pz, bz, bz2=None, survey_pz2=None, zeff=False, samplevar_lim=False,
               nkk=None):
    if nkk is None:
        nkk = Nkk(lensCl_interps, nolensCl_interps, Llls, terms=['TT', 'TE', 'EE', 'EB'],
                  thetab=thetab, DeltaT=DeltaT, iterative=iterative, pickleit=False)
    cgg = Cgg(Pk_interps, Llls, zmin, zmax, pz, bz, zeff=zeff, bz2=bz, survey_pz2=pz)
    ngg = Ngg(Llls, zmin, zmax, pz, nbar)
    ckg = Ckg(Pk_interps, Llls, zmin, zmax, pz, bz, zeff=zeff)
    ckk = Ckk(Pk_interps, Llls)
    return ({'gg': cgg, 'kg': ckg, 'gk': ckg, 'kk': ckk},
            {'gg': ngg, 'kg': np.zeros_like(ckg), 'gk': np.zeros_like(ckg), 'kk': nkk})


_FIELD = {'g': 0, 'k': 1}


def get_ClsCov(Llls, Cls, Nls, fsky, printit=False):
    """Gaussian covariance, eqn. (14) of arXiv:1710.09465; dict key 'abcd' -> Cov(C^{ab}, C^{cd}) per L (interloper.py:42-63)."""
    keys = list(Cls)
    L = np.asarray(Llls, dtype=np.float64)
    tot = np.empty((L.size, 2, 2))
    for a in 'gk':
        for b in 'gk':
            tot[:, _FIELD[a], _FIELD[b]] = np.asarray(Cls[a + b], dtype=np.float64) + np.asarray(Nls[a + b], dtype=np.float64)
    cov = _fisher.gaussian_cov(L, tot, [(_FIELD[k[0]], _FIELD[k[1]]) for k in keys], fsky)
    result = {}
    for i, key in enumerate(keys):
        for j, kkey in enumerate(keys):
            result[key + kkey] = cov[:, i, j].copy()
    if printit:
        for key in result:
            print(key, result[key])
    return result


def dp_DL(p, fid_sig8, fid_b1, DL):
    if p == 's8':
        return (2. / fid_sig8) * DL
    elif p == 'b1':
        return np.array([0.0, DL[1] / fid_b1, 2. * DL[2] / fid_b1])
    else:
        raise ValueError('\n\nDerivative with respect to given parameter is not available.\n\n')


def fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax, device=None, fused=True):
    """interloper.py:200-260: Fisher matrix of (sigma8, b1) from D = (Ckk, Ckg, Cgg) for L < LllMax and the parameter shift
    caused by the contaminated spectra xCls / xNls.  Returns dict(Fisher, iFisher, interim, dtheta, biased_sig8, biased_b1).
    `fused=False` forces the general route (get_ClsCov + fisher_accum) for inputs the one-launch route would also accept."""
    Llls = np.asarray(Llls)
    one = _fused_fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax, device) if fused else None
    if one is not None:
        Fisher, interim = one
        return _finish_bias(Fisher, interim, fid_sig8, fid_b1)
    covs = get_ClsCov(Llls, Cls, Nls, fsky)
    keep = Llls < LllMax                                                 # :221
    n = int(np.count_nonzero(keep))
    if not np.all(keep[:n]):
        raise ValueError('Llls must be increasing (the reference indexes the uncut arrays with the cut loop counter)')
    tr = ['kk', 'kg', 'gg']
    cov = np.empty((n, 3, 3))
    for i, a in enumerate(tr):
        for j, b in enumerate(tr):
            cov[:, i, j] = covs[a + b][:n]                               # :234-237
    DL = np.stack([np.asarray(Cls[t], dtype=np.float64)[:n] for t in tr], axis=1)          # :232
    dD = np.empty((n, 2, 3))
    dD[:, 0, :] = (2. / fid_sig8) * DL                                   # dp_DL('s8')
    dD[:, 1, 0] = 0.0
    dD[:, 1, 1] = DL[:, 1] / fid_b1
    dD[:, 1, 2] = 2. * DL[:, 2] / fid_b1
    dckg = np.asarray(Cls['kg'])[:n] - np.asarray(xCls['kg'])[:n]        # :224
    dcgg = (np.asarray(Cls['gg'])[:n] + np.asarray(Nls['gg'])[:n]) - (np.asarray(xCls['gg'])[:n] + np.asarray(xNls['gg'])[:n])
    DDL = np.stack([np.zeros(n), dckg, dcgg], axis=1)                    # :250
    w = np.asarray(nmodes, dtype=np.float64)[:n]
    Fisher, interim = _fisher.fisher_accum(cov, dD, w, dDelta=DDL, device=device)
    return _finish_bias(Fisher, interim, fid_sig8, fid_b1)


def _finish_bias(Fisher, interim, fid_sig8, fid_b1):
    iFisher = inv(Fisher)
    dtheta = np.zeros(2)
    for i in range(2):
        for j in range(2):
            dtheta[i] += iFisher[i, j] * interim[j]                      # interloper.py:255-258
    return {'Fisher': Fisher, 'iFisher': iFisher, 'interim': interim, 'dtheta': dtheta,
            'biased_sig8': fid_sig8 + dtheta[0], 'biased_b1': fid_b1 + dtheta[1]}


def _fused_fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax, device):
    """One launch (csrc/sweep.cu, a 1 x 1 design grid) when the inputs have the shape get_allCls gives them -- constant N_gg,
    N_kg = 0, C_gk = C_kg, increasing L: the per-multipole covariance, its solve and the Fisher / bias sums never leave the
    device.  Any other input takes the general get_ClsCov + fisher_accum route; the two agree to 1e-12
    (tests/test_sweep_gpu.py, tests/test_forecast_stage_gpu.py)."""
    from lbgcmb_b200.sweep import forecast_grid
    try:
        ngg, xngg = np.asarray(Nls['gg'], dtype=np.float64), np.asarray(xNls['gg'], dtype=np.float64)
        ok = (ngg.size > 0 and np.all(ngg == ngg.flat[0]) and np.all(xngg == xngg.flat[0])
              and not np.any(Nls['kg']) and not np.any(Nls['gk']) and not np.any(xNls['kg'])
              and (Cls['gk'] is Cls['kg'] or np.array_equal(Cls['gk'], Cls['kg']))
              and np.all(np.diff(Llls) > 0))
    except (KeyError, TypeError, ValueError):
        return None
    if not ok:
        return None
    out = forecast_grid(Llls, nmodes, Cls['kk'], Cls['gg'], Cls['kg'], ngg.flat[0], Nls['kk'], fsky, fid_sig8, fid_b1, LllMax,
                        xcgg=xCls['gg'], xckg=xCls['kg'], xngg=xngg.flat[0], device=device)
    return out[0, 0, 1:5].reshape(2, 2).copy(), out[0, 0, 5:7].copy()


def save_result(path, peakz, band, int_frac, res):
    """The dat/result4interloper_*.json layout of interloper.py:271-275 (`res` = the dict fisher_bias returns plus fid_sig8, fid_b1)."""
    import json
    output = {'peakz': peakz, 'fid_sig8': float(res['fid_sig8']), 'fid_b1': float(res['fid_b1']), 'biased_sig8': float(res['biased_sig8']),
              'biased_b1': float(res['biased_b1']), 'iFisher': np.asarray(res['iFisher']).tolist(), 'intfrac': int_frac}
    with open(path, 'w') as fp:
        json.dump(output, fp, sort_keys=True, indent=4)
    return output
