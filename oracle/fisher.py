"""CPU restatement of the S/N, covariance and Fisher stage (TEST INFRASTRUCTURE).

snr            snr.py:21-25
get_ClsCov     interloper.py:42-63        dp_DL  interloper.py:65-74
fisher_bias    interloper.py:200-260 (script body: per-L 3x3 covariance, inverse, Fisher and bias sums, final inverse)
mqw_fisher     mqw/py/mqw.py:18-127 given the shared C_gg (the function calls Cgg itself at :48)
rho_from_spectra   yu.py:86-108 (script body: shot noise on the diagonal, correlation matrix, inverse, rho_L)
Pinned against the unmodified reference by oracle/make_golden.py -> tests/golden/fisher.npz.
"""
from collections import OrderedDict

import numpy as np
from numpy.linalg import inv


def snr(Llls, Cls, Vars, nmodes, lmax=1.e5):
    result = (Cls[Llls <= lmax] / np.sqrt(Vars[Llls <= lmax])) ** 2.
    result = np.sum(nmodes[Llls <= lmax] * result)
    return np.sqrt(result)


def get_ClsCov(Llls, Cls, Nls, fsky):
    result = {}
    for key in Cls:
        for kkey in Cls:
            r = (Cls[key[0] + kkey[0]] + Nls[key[0] + kkey[0]]) * (Cls[key[1] + kkey[1]] + Nls[key[1] + kkey[1]])
            r = r + (Cls[key[0] + kkey[1]] + Nls[key[0] + kkey[1]]) * (Cls[key[1] + kkey[0]] + Nls[key[1] + kkey[0]])
            r = r / (2. * Llls + 1.)
            r = r / fsky
            result[key + kkey] = r
    return result


def dp_DL(p, fid_sig8, fid_b1, DL):
    if p == 's8':
        return (2. / fid_sig8) * DL
    elif p == 'b1':
        return np.array([0.0, DL[1] / fid_b1, 2. * DL[2] / fid_b1])
    raise ValueError('\n\nDerivative with respect to given parameter is not available.\n\n')


def fisher_bias(Llls, nmodes, Cls, Nls, xCls, xNls, fsky, fid_sig8, fid_b1, LllMax, get_ClsCov=get_ClsCov, dp_DL=dp_DL):
    """`get_ClsCov` / `dp_DL` can be swapped for the unmodified reference functions (oracle/make_golden.py)."""
    nparam = 2
    dtheta, interim = np.zeros(nparam), np.zeros(nparam)
    Fisher = np.zeros((nparam, nparam))
    covs = get_ClsCov(Llls, Cls, Nls, fsky)
    Cov_Lll = np.zeros((3, 3))
    Llls = Llls[Llls < LllMax]
    for LL, dummy in enumerate(Llls):
        dckg = Cls['kg'][LL] - xCls['kg'][LL]
        dcgg = (Cls['gg'][LL] + Nls['gg'][LL]) - (xCls['gg'][LL] + xNls['gg'][LL])
        DL = np.array([Cls['kk'][LL], Cls['kg'][LL], Cls['gg'][LL]])
        for i, tracer in enumerate(['kk', 'kg', 'gg']):
            for j, ttracer in enumerate(['kk', 'kg', 'gg']):
                Cov_Lll[i, j] = covs[tracer + ttracer][LL]
        iCov_Lll = inv(Cov_Lll)
        for i, pp in enumerate(['s8', 'b1']):
            for j, ss in enumerate(['s8', 'b1']):
                Fisher[i, j] += nmodes[LL] * np.dot(dp_DL(pp, fid_sig8, fid_b1, DL), np.dot(iCov_Lll, dp_DL(ss, fid_sig8, fid_b1, DL)))
        DDL = np.array([0.0, dckg, dcgg])
        iCovDDL = np.dot(iCov_Lll, DDL)
        for j, pparam in enumerate(['s8', 'b1']):
            interim[j] += nmodes[LL] * np.dot(dp_DL(pparam, fid_sig8, fid_b1, DL), iCovDDL)
    iFisher = inv(Fisher)
    for i in range(nparam):
        for j in range(nparam):
            dtheta[i] += iFisher[i, j] * interim[j]
    return {'Fisher': Fisher, 'iFisher': iFisher, 'interim': interim, 'dtheta': dtheta,
            'biased_sig8': fid_sig8 + dtheta[0], 'biased_b1': fid_b1 + dtheta[1]}


def mqw_fisher(cgg, Llls, bins, fsky):
    """bins[n, 7] = bp, bs, Np, Ns, wp, ws, wps per shell; returns (Fisher, Fii Schur-Limber, S)."""
    result = OrderedDict()
    for i, (bp, bs, Np, Ns, wp, ws, wps) in enumerate(np.asarray(bins)):
        r = {'Ns': Ns, 'Np': Np, 'num': (bp * bs * cgg) ** 2., 'denom': cgg * (bp * Np) ** 2. + wp}
        r['A0i'] = bp * Np * bs * Ns * cgg + wps
        r['Aii'] = (bs * Ns) ** 2. * cgg + ws
        r['A0i_j'] = bp * bs * Ns * cgg
        result[i] = r
    A00, sshift = 0.0, 0.0
    for z in result:
        A00 = A00 + result[z]['denom']
        sshift = sshift + result[z]['A0i'] ** 2. / result[z]['Aii']
    Fii = np.array([result[z]['Ns'] * fsky * np.sum((2. * Llls + 1) * (result[z]['num'] / A00)) for z in result])
    S = A00 / (A00 - sshift)
    for z in result:
        result[z]['r'] = result[z]['A0i'] / np.sqrt(A00 * result[z]['Aii'])
    n = len(result)
    Fisher = np.zeros((n, n))
    for i in result:
        Fisher[i, i] = fsky * np.sum((2. * Llls + 1) * S * result[i]['A0i_j'] * result[i]['A0i_j'] / result[i]['Aii'] / A00)
        for j in result:
            Fisher[i, j] += np.sum((2. * Llls + 1) * fsky * 2. * S ** 2. * result[i]['r'] * result[j]['r']
                                   * np.sqrt(1. / result[i]['Aii'] / result[j]['Aii']) * result[i]['A0i_j'] * result[j]['A0i_j'] / A00)
    return Fisher, Fii, S


def rho_from_spectra(kk, kg, gg, ngg):
    """One L: kk scalar, kg[nt], gg[nt, nt] (no noise), ngg[nt] shot noise -> rho_L (yu.py:86-108)."""
    kg = np.array(kg, dtype=np.float64)
    gg = np.array(gg, dtype=np.float64)
    for i in range(len(kg)):
        gg[i, i] += ngg[i]
    diag = np.diag(gg)
    rho = np.copy(gg)
    for i in np.arange(len(diag)):
        for j in np.arange(len(diag)):
            rho[i, j] /= np.sqrt(diag[i]) * np.sqrt(diag[j])
    kg /= (np.sqrt(kk) * np.sqrt(diag))
    irho = np.linalg.inv(rho)
    interim = 0.0
    for i in np.arange(len(kg)):
        for j in np.arange(len(kg)):
            interim += kg[i] * irho[i, j] * kg[j]
    return np.sqrt(interim)


def field_fisher(n_f, spectra, tot, dC, mode_weight):
    """Field form of the interloper.py:223-253 Fisher sum for n_f correlated Gaussian fields (no reference counterpart beyond n_f = 2):
    F[p, q] = sum_l w_l 1/2 Tr[C_l^-1 d_pC_l C_l^-1 d_qC_l], dense numpy.  With fields (kappa, g), spectra (kk, kg, gg) and
    w_l = nmodes (2l+1) fsky it equals fisher_bias()'s `dD^T Cov^-1 dD` sum (checked in tests/test_field_fisher_gpu.py)."""
    tot, dC = np.asarray(tot, dtype=np.float64), np.asarray(dC, dtype=np.float64)
    n_par, n_spec, n_ell = dC.shape
    F = np.zeros((n_par, n_par))
    Q = np.zeros((n_ell, n_par, n_par))
    for l in range(n_ell):
        C = np.zeros((n_f, n_f))
        D = np.zeros((n_par, n_f, n_f))
        for s, (a, b) in enumerate(spectra):
            C[a, b] = C[b, a] = tot[s, l]
            D[:, a, b] = dC[:, s, l]
            D[:, b, a] = dC[:, s, l]
        M = np.linalg.solve(C, D)                       # [n_par, n_f, n_f] = C^-1 d_pC
        Q[l] = 0.5 * np.einsum('pij,qji->pq', M, M)
        F += mode_weight[l] * Q[l]
    return F, Q
