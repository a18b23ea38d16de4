"""CPU restatement of the modi flavour background + Limber code (TEST INFRASTRUCTURE).

ModiCosmology   modi/cosmology.py:17-182 (Ea, Ha, chia by quad, D+ by scipy-1.2.1 romberg on 1000 ln a nodes, k_chil, pkalin,
                pkanlin = HALOFIT for z <= 4 else linear)
LensingX        modi/lensing.py:43-179 (kernels, weights, k meshes, cl_zeff, cl_zeff_bitracer, clz, cl_kkz; np.trapz over chi)
Pinned against the reference's own modi/cosmology.py and modi/lensing.py, loaded unmodified by oracle/refload.load_modi()
(tests/golden/modi_lensing.npz, oracle/make_golden.py)."""
import numpy as np
from scipy.integrate import quad
from scipy.interpolate import InterpolatedUnivariateSpline

from .halofit import halofit
from .romberg121 import romberg
from .shim.astropy.cosmology import FlatLambdaCDM


class ModiCosmology(object):
    def __init__(self, M=0.292, L=None, H0=100., B=None, sig8=0.819, klin=None, plin=None):
        self.M, self.L, self.B, self.H0, self.sig8 = M, (1 - M if L is None else L), B, H0, sig8
        self.cin = 1000. / (299792458.)
        self.cosmo = FlatLambdaCDM(H0=H0, Om0=M, Ob0=B, name="cosmo")
        self.klin, self.plin = klin, plin
        logx = np.linspace(-20., 0, 1000)
        x = np.exp(logx)
        kernel = lambda loga: (np.exp(loga) * self.Ea(a=np.exp(loga))) ** -3. * np.exp(loga)
        y = self.Ea(a=x) * np.array([romberg(kernel, logx.min(), loga, vec_func=True) for loga in logx])
        self.Dplus = InterpolatedUnivariateSpline(x, y)

    def _za(self, z, a):
        if a is None:
            return z, 1. / (1. + z)
        return 1. / a - 1., a

    def Ea(self, z=None, a=None):
        z, a = self._za(z, a)
        return a ** -1.5 * (self.M + (1. - self.M - self.L) * a + self.L * a ** 3.) ** 0.5

    def Ha(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.H0 * self.Ea(a=a)

    def Dgrow(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.Dplus(a) / self.Dplus(1.)

    def chia(self, z=None, a=None):
        fac = 100. * self.cin
        f = lambda x: (self.Ea(a=x) * x ** 2) ** -1.
        z, a = self._za(z, a)
        if type(a) == np.ndarray:
            return np.array([quad(f, ai, 1)[0] for ai in a]) / fac
        return quad(f, a, 1)[0] / fac

    def k_chil(self, l, z=None, a=None):
        z, a = self._za(z, a)
        return (l + 0.5) / self.chia(a=a)

    def pkalin(self, z=None, a=None):
        z, a = self._za(z, a)
        return self.klin, self.plin * self.Dgrow(a=a) ** 2.

    def pkanlin(self, z=None, a=None, sig8=None):
        z, a = self._za(z, a)
        sig8 = self.sig8 if sig8 is None else sig8
        if z > 4:
            return self.pkalin(z, a)
        pdiml = self.plin * self.klin ** 3 / (2. * np.pi ** 2)
        pgive = pdiml * self.Dgrow(a=a) ** 2.
        pndim = halofit(k=self.klin, delta_k=pgive, sigma_8=sig8, z=1. / a - 1., cosmo=self.cosmo, takahashi=True)
        return self.klin, pndim * (2. * np.pi ** 2) / self.klin ** 3


class LensingX(object):
    def __init__(self, z, dndz, cosmo, l=None, dz=0.5, nz=100, ggnorm=None):
        self.l = np.logspace(1, 3.47, 500) if l is None else l
        self.z, self.dz, self.nz, self.kknz, self.dndz, self.cosmo, self.ggnorm = z, dz, nz, 20 * nz, dndz, cosmo, ggnorm
        self.zs = np.linspace(z - dz / 2., z + dz / 2., nz)
        self.kkzs = np.linspace(0.001, 1100., self.kknz)
        self.chis = cosmo.chia(z=self.zs)
        self.kerg = self.kernel_g(self.zs, dndz)
        self.norm = np.trapezoid(self.kerg, self.chis)
        if ggnorm is not None:
            self.norm = ggnorm
        self.kerg = self.kerg / self.norm
        self.kercmb = self.kernel_cmb(z=self.zs)
        self.ggwts = self.kerg ** 2. / self.chis ** 2.
        self.kgwts = self.kercmb * self.kerg / self.chis ** 2.
        self.kkwts = self.kernel_cmb(z=self.kkzs) * self.kernel_cmb(z=self.kkzs) / cosmo.chia(z=self.kkzs) ** 2.
        self.dggwtsdN = 2. * self.kerg / self.chis ** 2. * cosmo.Ha(z=self.zs) / self.norm
        self.dkgwtsdN = self.kercmb / self.chis ** 2. * cosmo.Ha(z=self.zs) / self.norm
        self.kmesh = np.zeros([self.l.size, self.zs.size])
        self.kk_kmesh = np.zeros([self.l.size, self.kkzs.size])
        for i in range(self.zs.size):
            self.kmesh[:, i] = cosmo.k_chil(self.l, z=self.zs[i])
        for i in range(self.kkzs.size):
            self.kk_kmesh[:, i] = cosmo.k_chil(self.l, z=self.kkzs[i])

    def kernel_cmb(self, z=None, a=None):
        z, a = self.cosmo._za(z, a)
        chistar = self.cosmo.chia(z=1100.)
        chi = self.cosmo.chia(z)
        return chi * (chistar - chi) / (chistar) * (3. * self.cosmo.M * self.cosmo.H0 ** 2. / (2. * a)) * self.cosmo.cin ** 2.

    def kernel_g(self, z, dndz):
        return self.cosmo.Ha(z=z) * dndz(z)

    def cl_zeff(self, kp, auto=False):
        pval = np.interp(self.kmesh, *kp)
        return np.trapezoid(pval * (self.ggwts if auto else self.kgwts), self.chis, axis=-1)

    def cl_zeff_bitracer(self, kp, dndz2):
        pval = np.interp(self.kmesh, *kp)
        kerg2 = self.kernel_g(self.zs, dndz2)
        kerg2 = kerg2 / np.trapezoid(kerg2, self.chis)
        return np.trapezoid(pval * (self.kerg * kerg2 / self.chis ** 2.), self.chis, axis=-1)

    def clz(self, ikpz, auto=False):
        pval = np.zeros_like(self.kmesh)
        for i in range(self.nz):
            pval[:, i] = np.interp(self.kmesh[:, i], *ikpz(z=self.zs[i]))
        return np.trapezoid(pval * (self.ggwts if auto else self.kgwts), self.chis, axis=-1)

    def cl_kkz(self, ikpz):
        pval = np.zeros_like(self.kk_kmesh)
        for i in range(self.kknz):
            pval[:, i] = np.interp(self.kk_kmesh[:, i], *ikpz(z=self.kkzs[i]), left=0.0, right=0.0)
        return np.trapezoid(pval * self.kkwts, self.cosmo.chia(z=self.kkzs), axis=-1)
