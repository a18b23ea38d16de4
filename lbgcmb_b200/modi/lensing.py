"""Lensing cross-spectra in the thin-slice / Limber approximation, modi flavour; mirrors reference
modi/lensing.py:14-179 (`prefactor`, `CosmoPar`, `LensingX`).

Host (as in the reference, <= 2000 doubles): the z grids, chi(z) by quad, the kernels W_g, W_kappa and the weight vectors
(modi/lensing.py:67-96).  Device (csrc/limber.cu): P(k, z) on every z node (growth + HALOFIT), np.interp onto the
k = (l + 1/2)/chi mesh and the trapezoid over chi for every l (`cl_zeff`, `cl_zeff_bitracer`, `clz`, `cl_kkz`).
The (l x z) k meshes of modi/lensing.py:99-107 are built lazily: the kernels never need them."""
import numpy as np

from lbgcmb_b200 import _ffi
from lbgcmb_b200.modi import cosmology
from lbgcmb_b200.quadrature import trapz_weights_x


def prefactor(ells, n=1):
    if n == 0:
        return 1.0
    if n == 1:
        return (2. * ells + 1.0) / (4. * np.pi)
    if n == 2:
        return ells * (ells + 1.0) / (2. * np.pi)


class CosmoPar:
    M = 0.292
    L = 1 - M
    H0 = 100.
    h = 0.69


class PkZ(object):
    """`ikpz`: callable (z=...) -> [k, P(k, z) b(z)^power], the shape of the lambdas in modi/lensing.py:214-218
    (`mmz`, `hmz`, `hhz`).  LensingX recognises it and keeps the whole projection on the device."""

    def __init__(self, cosmo, nonlinear=True, bias=None, power=1, sig8=None):
        self.cosmo, self.nonlinear, self.bias, self.power, self.sig8 = cosmo, nonlinear, bias, power, sig8

    def amplitude(self, zs):
        if self.bias is None:
            return np.ones(len(zs))
        return np.array([float(self.bias(z)) for z in zs]) ** self.power

    def __call__(self, z):
        k, p = self.cosmo.pkanlin(z=z, sig8=self.sig8) if self.nonlinear else self.cosmo.pkalin(z=z)
        return [k, p * self.amplitude([z])[0]]


class LensingX:
    ''' Lensing cross-spectra in the thin-slice approximation.'''

    def __init__(self, z, dndz, cosmo=None, l=None, dz=0.5, nz=100, ggnorm=None):
        if cosmo is None:
            raise ValueError('pass a modi.cosmology.Cosmology: the reference default reads a file under /Users/M.J.Wilson '
                             '(modi/lensing.py:30,46)')
        if l is None:
            l = np.logspace(1, 3.47, 500)
        if dndz is None:
            dndz = lambda x: 1
        self.ggnorm = ggnorm
        self.z = z
        self.l = l
        self.dz = dz
        self.nz = nz
        self.kknz = 20 * self.nz
        self.dndz = dndz
        self.cosmo = cosmo
        self._clsetup()

    def _clsetup(self):
        zmin = self.z - self.dz / 2.
        zmax = self.z + self.dz / 2.
        self.zs = np.linspace(zmin, zmax, self.nz)
        self.kkzs = np.linspace(0.001, 1100., self.kknz)
        self.chis = self.cosmo.chia(z=self.zs)
        self.kkchis = self.cosmo.chia(z=self.kkzs)
        self.kerg = self.kernel_g(z=self.zs, dndz=self.dndz)
        self.norm = np.trapezoid(self.kerg, self.chis)
        if self.ggnorm is not None:
            print("External gg-norm given = %0.3f, gg-norm estimated from dndz = %0.3f; Overriding estimated norm. Be wary!!!"
                  % (self.ggnorm, self.norm))
            self.norm = self.ggnorm
        self.kerg = self.kerg / self.norm
        self.kercmb = self.kernel_cmb(z=self.zs)
        self.ggwts = self.kerg ** 2. / self.chis ** 2.
        self.kgwts = self.kercmb * self.kerg / self.chis ** 2.
        kk = self.kernel_cmb(z=self.kkzs)
        self.kkwts = kk * kk / self.kkchis ** 2.
        self.dggwtsdN = 2. * self.kerg / self.chis ** 2. * self.cosmo.Ha(z=self.zs) / self.norm
        self.dkgwtsdN = self.kercmb / self.chis ** 2. * self.cosmo.Ha(z=self.zs) / self.norm
        self._tw = trapz_weights_x(self.chis)             # np.trapz(cl, self.chis, axis=-1)
        self._kktw = trapz_weights_x(self.kkchis)

    @property
    def kmesh(self):
        return (self.l[:, None] + 0.5) / self.chis[None, :]

    @property
    def kk_kmesh(self):
        return (self.l[:, None] + 0.5) / self.kkchis[None, :]

    def kernel_cmb(self, z=None, a=None):
        z, a = self.cosmo._za(z, a)
        chistar = self.cosmo.chia(z=1100.)
        chi = self.cosmo.chia(z)
        f = chi * (chistar - chi) / (chistar)
        fac = 3. * self.cosmo.M * self.cosmo.H0 ** 2. / (2. * a)
        return f * fac * self.cosmo.cin ** 2.

    def kernel_g(self, z, dndz):
        Ha = self.cosmo.Ha(z=z)
        return Ha * dndz(z)

    # ---- device projections ---------------------------------------------------------------------------------------
    def _project_table(self, kp, chis, weights, oob):
        """sum_z weights[z] np.interp((l + 1/2)/chi_z, k, P) for one caller-supplied (k, P) table."""
        k, P = kp
        return self.cosmo.limber_rows(self.l, chis, weights[None, :], k, np.asarray(P)[None, :], np.zeros(len(chis), dtype=np.int32), oob)[0]

    def cl_zeff(self, kp, auto=False):
        ''' Assumes the power spectrum is constant across the redshift bin '''
        wts = self.ggwts if auto else self.kgwts
        return self._project_table(kp, self.chis, wts * self._tw, oob=1)

    def cl_zeff_bitracer(self, kp, dndz2):
        ''' Cross-spectra with two galaxy tracers; assumes constant power spectrum across redshift bin '''
        kerg2 = self.kernel_g(z=self.zs, dndz=dndz2)
        kerg2 = kerg2 / np.trapezoid(kerg2, self.chis)
        ggwts2 = self.kerg * kerg2 / self.chis ** 2.
        return self._project_table(kp, self.chis, ggwts2 * self._tw, oob=1)

    def _project_z(self, ikpz, zs, chis, weights, oob, keep=None):
        if isinstance(ikpz, PkZ):
            w = weights * ikpz.amplitude(zs)
            out = ikpz.cosmo.limber(self.l, zs, chis, w[None, :], nonlinear=ikpz.nonlinear, oob=oob, sig8=ikpz.sig8, keep=keep)
            return None if keep is not None else out[0]
        if keep is not None:
            raise ValueError('keep= needs a PkZ table (device P(k, z))')
        tabs = [ikpz(z=z) for z in zs]                     # a foreign callable is the caller's input: sampled on the host
        k = np.asarray(tabs[0][0], dtype=np.float64)
        for t in tabs:
            if len(t[0]) != k.size or not np.array_equal(np.asarray(t[0], dtype=np.float64), k):
                raise ValueError('ikpz must return every P(k, z) on one k grid')
        rows = np.array([t[1] for t in tabs], dtype=np.float64)
        return self.cosmo.limber_rows(self.l, chis, weights[None, :], k, rows, np.arange(len(zs), dtype=np.int32), oob)[0]

    def clz(self, ikpz, auto=False):
        ''' Assuming interpolating function for Pk(z) across the redshift bin '''
        wts = self.ggwts if auto else self.kgwts
        return self._project_z(ikpz, self.zs, self.chis, wts * self._tw, oob=1)

    def cl_kkz(self, ikpz, keep=None):
        ''' Assuming interpolating function for Pk(z) across the redshift bin.  keep = (set, set_rows, row): left on the device '''
        if keep is not None:
            keep = (keep[0], keep[1], [keep[2]])
        return self._project_z(ikpz, self.kkzs, self.kkchis, self.kkwts * self._kktw, oob=0, keep=keep)
