"""Tomographic CMB-lensing x galaxy Fisher forecast over cosmological and bias parameters (BASELINE configs[3]: kappa + 6 tracers x 20
redshift bins, parameters sigma8 / sum m_nu / Omega_m and one bias amplitude per tracer).

The reference has no single driver for it: tracers as in yu.py:44-48, bins of mqw/py/mqw.py:20, Limber projections in
the modi flavour (modi/lensing.py:69-179: W_g = H dN/dz normalised by its trapezoid over chi, trapezoid over chi of W W / chi^2 P), the
Fisher sum of interloper.py:200-253 taken over all fields at once (fisher.field_fisher), and its only cosmological derivative is the
analytic sigma8 one of interloper.py:65-74.  Here every cosmology of the finite-difference stencil is built on the device: linear table
from the Eisenstein-Hu fit with massive neutrinos (lbg_eh_linear_table, EisensteinAndHu.c:102-264), growth tables for the whole stencil
in one launch (lbg_growth_table), HALOFIT rows and the Limber sums (csrc/limber.cu), then one lbg_field_fisher launch.  With the
multipoles sharded over ranks the partial matrices meet in one all-reduce (dist.allreduce_sum)."""
import numpy as np

from lbgcmb_b200 import dist, eisenstein_hu, fisher
from lbgcmb_b200.modi.cosmology import Cosmology, growth_tables
from lbgcmb_b200.modi.lensing import LensingX, PkZ
from lbgcmb_b200.quadrature import trapz_weights_x

FIDUCIAL = {'sig8': 0.819, 'M': 0.292, 'B': 0.0468, 'h': 0.69, 'ns': 0.965, 'mnu': 0.06}       # M, h: modi/lensing.py:24-28
STEPS = {'sig8': 0.01, 'M': 0.005, 'B': 0.001, 'h': 0.01, 'ns': 0.01, 'mnu': 0.02}                 # absolute, central differences
K_GRID = np.logspace(-4., 2., 1200)


_GL_X, _GL_W = np.polynomial.legendre.leggauss(48)


class Background(Cosmology):
    """modi Cosmology whose comoving distance is a 48-point Gauss-Legendre rule in u = sqrt(a), vectorised over redshift, in place of one
    scipy `quad` per redshift (modi/cosmology.py:134-150): chi = (c / 100) int_a^1 da / (a^2 E) = (c / 100) int 2 du / sqrt(M + (1 - M - L) u^2
    + L u^6), smooth down to a = 0.  Same value to 1e-13 (tests/test_tomography_cpu.py); a forecast stencil needs ~7000 distances per
    background."""

    def chia(self, z=None, a=None):
        z, a = self._za(z, a)
        u0 = np.sqrt(np.atleast_1d(np.asarray(a, dtype=np.float64)))
        half = 0.5 * (1. - u0)
        u = u0[:, None] + half[:, None] * (_GL_X[None, :] + 1.)
        f = 2. / np.sqrt(self.M + (1. - self.M - self.L) * u ** 2. + self.L * u ** 6.)
        out = half * np.dot(f, _GL_W) / (100. * self.cin)
        return out if np.ndim(a) else float(out[0])


def omega_hdm(mnu, h):
    """Density parameter of sum m_nu [eV] in massive neutrinos: m / (93.14 h^2 eV)."""
    return mnu / (93.14 * h * h)


def make_cosmology(theta, Dplus_table='device', k=None, device=None):
    """modi Cosmology for the parameter dict `theta` (keys of FIDUCIAL) with its linear table built on the device by the E-H fit,
    sigma8-normalised (so sigma8 is a parameter in its own right and the other ones move the shape at fixed sigma8).
    sum m_nu enters through the z = 0 transfer-function shape (T_cb+nu) and HALOFIT's f_nu terms; the growth factor stays the
    scale-independent LCDM one ("sum m_nu limitation")."""
    k = K_GRID if k is None else np.asarray(k, dtype=np.float64)
    M, h = theta['M'], theta['h']
    plin = eisenstein_hu.linear_power(k, M, theta['B'], 1. - M, h, theta['ns'], theta['sig8'], omega_hdm=omega_hdm(theta['mnu'], h),
                                      degen_hdm=1, device=device)
    c = Background(M=M, B=theta['B'], sig8=theta['sig8'], klin=k, plin=plin, device=device, Dplus_table=Dplus_table)
    c.cosmo.Onu0 = omega_hdm(theta['mnu'], h)        # f_nu = Onu0 / Om0 of HALOFIT's neutrino terms (modi/halofit.py:149,164,204-206)
    return c


def field_spectra(cosmo, ell, tracers, nodes_per_bin=10, nonlinear=True, kk=True, geometry=None, keep_set=None):
    """C_kk [n_ell] (None without `kk`), C_kg [n_field, n_ell], C_gg [n_field, n_ell] for every redshift bin of every tracer; ONE batched
    Limber launch for all tracers (block-diagonal weights over the tracers' concatenated z grids).  tracers: dicts with 'edges' [n_bin + 1] (contiguous bins),
    'dndz' and 'bz' callables.  Bin i integrates over its own `nodes_per_bin` + 1 nodes of the tracer's z grid (edge nodes belong to both
    neighbours), normalised over the bin as modi/lensing.py:78-87.  `geometry`: a dict the caller keeps between calls with the SAME
    ell, tracers and nodes_per_bin; the distances and weight vectors (host `quad`s, modi/cosmology.py:134-150) depend on (M, L) only and
    are then computed once per background instead of once per cosmology.
    keep_set = s: nothing comes back to the host; spectra set s on the device holds [kk | kg rows | gg rows] (2 n_field + 1 rows,
    the spectrum order tomographic_fisher uses) for fisher.field_fisher_sets, and the return value is None."""
    ell = np.asarray(ell, dtype=np.float64)
    key = (cosmo.M, cosmo.L, cosmo.H0)
    geo = None if geometry is None else geometry.get(key)
    if geo is None:
        lx = LensingX(z=1.0, dndz=None, cosmo=cosmo, l=ell)
        m = int(nodes_per_bin)
        per_tracer = []
        for t in tracers:
            edges = np.asarray(t['edges'], dtype=np.float64)
            nb = edges.size - 1
            zs = np.concatenate([np.linspace(edges[i], edges[i + 1], m + 1)[:-1] for i in range(nb)] + [edges[-1:]])
            chis = cosmo.chia(z=zs)
            ker = cosmo.Ha(z=zs) * t['dndz'](zs)                                  # modi/lensing.py:109-112
            kcmb = lx.kernel_cmb(z=zs)
            b = np.array([float(t['bz'](z)) for z in zs])
            w = np.zeros((2 * nb, zs.size))
            for i in range(nb):
                s = slice(i * m, (i + 1) * m + 1)
                tw = trapz_weights_x(chis[s])
                kg_i = ker[s] / np.dot(tw, ker[s])                                # normalised over the bin, :78-87
                w[i, s] = kcmb[s] * kg_i / chis[s] ** 2. * b[s] * tw              # kgwts, :90
                w[nb + i, s] = kg_i ** 2. / chis[s] ** 2. * b[s] ** 2. * tw       # ggwts, :89
            per_tracer.append((nb, zs, chis, w))
        # ONE batched Limber call for all tracers: their z nodes one after another, the weights block-diagonal (a batch block of 8
        # entries holds bins of one tracer -- 2 n_bin is a multiple of 8 for the 20-bin case -- and the kernel skips the z nodes none of
        # a block's entries weights, so the zeros cost nothing); one P(k, z) table build per cosmology instead of one per tracer
        nz_all = sum(zs.size for _, zs, _, _ in per_tracer)
        nb_all = sum(nb for nb, _, _, _ in per_tracer)
        z_all = np.concatenate([zs for _, zs, _, _ in per_tracer])
        chi_all = np.concatenate([chis for _, _, chis, _ in per_tracer])
        w_all = np.zeros((2 * nb_all, nz_all))
        rows_kg, rows_gg, r0, c0 = [], [], 0, 0
        for nb, zs, chis, w in per_tracer:
            w_all[r0:r0 + 2 * nb, c0:c0 + zs.size] = w
            rows_kg.append(np.arange(r0, r0 + nb))
            rows_gg.append(np.arange(r0 + nb, r0 + 2 * nb))
            r0 += 2 * nb
            c0 += zs.size
        geo = (lx, (z_all, chi_all, w_all, np.concatenate(rows_kg), np.concatenate(rows_gg)))
        if geometry is not None:
            geometry[key] = geo
    lx, (z_all, chi_all, w_all, rows_kg, rows_gg) = geo
    if keep_set is not None:
        n_fld = rows_kg.size
        dest = np.empty(2 * n_fld, dtype=np.int32)
        dest[rows_kg] = 1 + np.arange(n_fld)
        dest[rows_gg] = 1 + n_fld + np.arange(n_fld)
        lx.cl_kkz(PkZ(cosmo, nonlinear=nonlinear), keep=(keep_set, 2 * n_fld + 1, 0))
        cosmo.limber(ell, z_all, chi_all, w_all, nonlinear=nonlinear, oob=1, keep=(keep_set, 2 * n_fld + 1, dest))
        return None
    ckk = lx.cl_kkz(PkZ(cosmo, nonlinear=nonlinear)) if kk else None
    out = cosmo.limber(ell, z_all, chi_all, w_all, nonlinear=nonlinear, oob=1)
    return ckk, out[rows_kg], out[rows_gg]



def tomographic_fisher(tracers, ell, nmodes, fsky, nkk, fid=None, cosmo_params=('sig8', 'mnu'), steps=None, nodes_per_bin=10,
                       nonlinear=True, k=None, device=None, sharded=False, spectra_fn=None, accum=None, return_spectra=True, host_arrays=False, geometry=None):
    """Fisher matrix over `cosmo_params` + one bias amplitude per tracer (b(z) -> A_t b(z), fiducial A_t = 1) from kappa and every
    galaxy bin: fields (kappa, g_1 .. g_n), spectra kk, kappa g_i, g_i g_i (bins of one tracer are disjoint; cross-tracer overlaps
    are neglected), noise N_kk = `nkk` [n_ell] and 1 / nbar_i on the galaxy autos (tracers[t]['nbar']: per bin, sr^-1).

    Cosmological derivatives are central differences over cosmologies built on the device; bias derivatives are analytic
    (C_kg / A, 2 C_gg / A).  n_par <= 8.  sharded=True: this rank takes ell[rank::world] and the matrices are all-reduced.
    Returns dict: params, F, cov (= F^-1), sigma (marginalised 1-sigma), spectra (fiducial kk, kg, gg on this rank's multipoles).
    `spectra_fn(theta, ell) -> (ckk, ckg, cgg)` and `accum` (signature of fisher.field_fisher) replace the device calls in the CPU
    (gloo) tests of the sharding logic; the product path leaves both None and then keeps every spectrum on the device (spectra sets:
    Cosmology.limber(keep=...), fisher.field_fisher_sets); host_arrays=True selects the host-array route of the same device kernels
    (spectra copied back, sums and differences in numpy, fisher.field_fisher_amp), kept for the parity test between the two.
    `geometry`: a dict the caller keeps between calls with the SAME ell, tracers and nodes_per_bin (a loop over noise curves or sky
    fractions): distances and Limber weight vectors are then built once per background instead of once per call.  return_spectra=False skips the read-back of the fiducial spectra."""
    fid = dict(FIDUCIAL, **(fid or {}))
    steps = dict(STEPS, **(steps or {}))
    ell, nmodes, nkk = (np.asarray(a, dtype=np.float64) for a in (ell, nmodes, nkk))
    if sharded:
        idx = dist.shard_indices(ell.size)
        ell, nmodes, nkk = ell[idx], nmodes[idx], nkk[idx]
    names = list(cosmo_params) + ['b_%s' % t.get('name', i) for i, t in enumerate(tracers)]
    if len(names) > 8:
        raise ValueError('at most 8 parameters (lbg_field_fisher): %d cosmological + %d tracers' % (len(cosmo_params), len(tracers)))
    for p in cosmo_params:
        if p not in FIDUCIAL:
            raise ValueError('unknown cosmological parameter %r (one of %s)' % (p, sorted(FIDUCIAL)))
    thetas = [fid]
    for p in cosmo_params:
        thetas += [dict(fid, **{p: fid[p] + steps[p]}), dict(fid, **{p: fid[p] - steps[p]})]
    nb = [len(t['edges']) - 1 for t in tracers]
    nf = 1 + sum(nb)
    ngg = np.concatenate([np.broadcast_to(1. / np.asarray(t['nbar'], dtype=np.float64), (n,)) for t, n in zip(tracers, nb)])
    spectra = [(0, 0)] + [(0, i) for i in range(1, nf)] + [(i, i) for i in range(1, nf)]
    n_c = len(cosmo_params)
    # bias amplitudes: d C_kg / d A = C_kg, d C_gg / d A = 2 C_gg on the tracer's own bins (the analytic dp_DL('b1') of
    # interloper.py:65-74): described by a (tracer x spectrum) factor matrix, their rows are formed on the device
    amp = np.zeros((len(tracers), len(spectra)))
    o = 0
    for t, n in enumerate(nb):
        amp[t, 1 + o:1 + o + n] = 1.0
        amp[t, nf + o:nf + o + n] = 2.0
        o += n
    w = nmodes * (2. * ell + 1.) * fsky
    if spectra_fn is None and accum is None and not host_arrays:
        # product path: every spectrum stays on the device.  Cosmology i of the stencil fills spectra set i; the noise is set
        # len(thetas); sums, differences and amplitude rows are formed there (lbg_field_fisher_sets)
        growth = growth_tables([t['M'] for t in thetas], device=device)       # the whole stencil in one launch
        geometry = {} if geometry is None else geometry
        for i, (theta, y) in enumerate(zip(thetas, growth)):
            c = make_cosmology(theta, Dplus_table=y, k=k, device=device)
            field_spectra(c, ell, tracers, nodes_per_bin, nonlinear, geometry=geometry, keep_set=i)
        noise_set = len(thetas)
        fisher.spectra_set_fill(noise_set, np.concatenate([np.zeros(nf), ngg]), ell.size, curves={0: nkk}, device=device)
        diffs = [(1 + 2 * j, 2 + 2 * j, 2. * steps[p]) for j, p in enumerate(cosmo_params)]
        F = fisher.field_fisher_sets(nf, spectra, ell.size, 0, noise_set, diffs, amp, w, device=device)
        if return_spectra:
            fidu = fisher.spectra_set_get(0, 0, 2 * nf - 1, ell.size, device=device)
            ckk, ckg, cgg = fidu[0], fidu[1:nf], fidu[nf:]
        else:
            ckk = ckg = cgg = None
    else:
        if spectra_fn is None:
            growth = growth_tables([t['M'] for t in thetas], device=device)
            spec, geometry = [], {}
            for theta, y in zip(thetas, growth):
                c = make_cosmology(theta, Dplus_table=y, k=k, device=device)
                spec.append(field_spectra(c, ell, tracers, nodes_per_bin, nonlinear, geometry=geometry))
        else:
            spec = [spectra_fn(theta, ell) for theta in thetas]
        ckk, ckg, cgg = spec[0]
        tot = np.concatenate([(ckk + nkk)[None], ckg, cgg + ngg[:, None]])
        dC = np.empty((n_c, len(spectra), ell.size))            # explicit rows for the cosmological parameters only
        for j, p in enumerate(cosmo_params):
            up, dn = spec[1 + 2 * j], spec[2 + 2 * j]
            for sl, x, y in ((slice(0, 1), up[0], dn[0]), (slice(1, nf), up[1], dn[1]), (slice(nf, 2 * nf - 1), up[2], dn[2])):
                np.subtract(x, y, out=dC[j, sl].reshape(np.shape(x)))
                dC[j, sl] /= 2. * steps[p]
        sig = np.concatenate([ckk[None], ckg, cgg])
        if accum is None:
            F = fisher.field_fisher_amp(nf, spectra, tot, dC, amp, sig, w, device=device)
        else:
            full = np.concatenate([dC, amp[:, :, None] * sig[None]])
            F = accum(nf, spectra, tot, full, w)
    if sharded:
        F, = dist.allreduce_sum(F)
    cov = np.linalg.inv(F)
    return {'params': names, 'F': F, 'cov': cov, 'sigma': np.sqrt(np.diag(cov)), 'fiducial': fid,
            'spectra': {'kk': ckk, 'kg': ckg, 'gg': cgg, 'ngg': ngg}, 'fields': nf, 'n_cosmologies': len(thetas)}


# ---- f sigma8(z) as a derived quantity ---------------------------------------------------------------------------------------
# BASELINE configs[3] lists "b(z) / f sigma8 / sum m_nu".  Angular kappa-g / g-g spectra in the Limber approximation carry no
# redshift-space-distortion term, so f sigma8 is not a parameter these observables constrain on their own; what the forecast can state
# is the error on f(z) sigma8(z) INSIDE the model (flat LCDM + the parameters varied), propagated from the Fisher covariance.
def growth_rate(cosmo, z):
    """f = d ln D / d ln a.  The growth table is D ~ y = E(a) int_0^a da' / (a' E)^3 (modi/cosmology.py:92-111), whose logarithmic
    derivative is closed-form given y: f = d ln E / d ln a + 1 / (a^2 E^2 y)."""
    a = 1. / (1. + np.asarray(z, dtype=np.float64))
    M, L = cosmo.M, cosmo.L
    K = 1. - M - L
    E2 = M * a ** -3. + K * a ** -2. + L
    dlnE = (-3. * M * a ** -3. - 2. * K * a ** -2.) / (2. * E2)
    return dlnE + 1. / (a ** 2. * E2 * cosmo.Dplus(a))


def fsigma8(cosmo, z):
    """f(z) sigma8 D(z) / D(0)."""
    return growth_rate(cosmo, z) * cosmo.sig8 * cosmo.Dgrow(z=np.asarray(z, dtype=np.float64))


def fsigma8_forecast(result, zs, steps=None, device=None, growth='device'):
    """1-sigma errors on f sigma8 at redshifts `zs` from a tomographic_fisher result: J cov J^T with J = d (f sigma8)(z) / d theta by
    central differences over the cosmological parameters of the result (bias amplitudes do not enter; of the FIDUCIAL keys only sig8 and
    M move the scale-independent growth modelled here -- sum m_nu changes the transfer-function shape only).  The
    stencil's growth tables come from one lbg_growth_table launch (growth='host': the reference's Romberg route, CPU tests)."""
    zs = np.atleast_1d(np.asarray(zs, dtype=np.float64))
    steps = dict(STEPS, **(steps or {}))
    fid = result['fiducial']
    cosmo_params = [p for p in result['params'] if p in FIDUCIAL]
    thetas = [fid]
    for p in cosmo_params:
        thetas += [dict(fid, **{p: fid[p] + steps[p]}), dict(fid, **{p: fid[p] - steps[p]})]
    tabs = growth_tables([t['M'] for t in thetas], device=device) if growth == 'device' else [None] * len(thetas)
    vals = [fsigma8(Background(M=t['M'], B=t['B'], sig8=t['sig8'], Dplus_table=y, device=device), zs) for t, y in zip(thetas, tabs)]
    J = np.zeros((zs.size, len(result['params'])))
    for j, p in enumerate(cosmo_params):
        J[:, result['params'].index(p)] = (vals[1 + 2 * j] - vals[2 + 2 * j]) / (2. * steps[p])
    cov = J @ result['cov'] @ J.T
    return {'z': zs, 'fsigma8': vals[0], 'sigma': np.sqrt(np.diag(cov)), 'cov': cov, 'jacobian': J}
