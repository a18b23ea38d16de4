"""Instrumental and reconstruction ("Fisher") noise in the modi form; mirrors reference modi/nkk.py:22-261
(`N_inst`, `N_fisher`, `N_fisherTE`, `CMB`).

The (l x phi) grids of modi/nkk.py:40-52 are never materialised: the nodes and scipy-1.2.1 Simpson weights are built here
exactly as the reference builds them and the integrand runs in csrc/n0_modi.cu (lbg_n0_modi).  The spline arguments are
table-backed `interpolate` objects (unit-spaced knots, as CLASS multipoles are); there is no CPU path."""
import numpy

from lbgcmb_b200 import _ffi
from lbgcmb_b200.DetectorNoise import DetectorNoise
from lbgcmb_b200.quadrature import simps_weights_x

MODE_BITS = {'TT': 1, 'TE': 2, 'EE': 4, 'EB': 8}
last_stats = {}


class interpolate(object):
    """InterpolatedUnivariateSpline(x, y) (k = 3, extrapolating) for unit-spaced x, evaluated on the GPU."""

    def __init__(self, x, y):
        x = numpy.asarray(x, dtype=numpy.float64)
        if x.ndim != 1 or x.size < 4 or numpy.max(numpy.abs(numpy.diff(x) - 1.0)) > 1e-9:
            raise ValueError('interpolate needs >= 4 unit-spaced knots (CLASS multipoles); general knot vectors are not supported')
        self.x0 = float(x[0])
        self.x = x
        self.y = _ffi.f64(y)
        if self.y.shape != x.shape:
            raise ValueError('x and y differ in shape')

    def __call__(self, x):
        xa = _ffi.f64(numpy.atleast_1d(x))
        out = numpy.empty(xa.size)
        ctx = _ffi.context()
        _ffi.check(_ffi.lib().lbg_eval_spline_ext(ctx.handle, self.y.size, self.x0, _ffi._dp(self.y), xa.size,
                                                  _ffi._dp(xa.ravel()), _ffi._dp(out)))
        out = out.reshape(xa.shape)
        return out if numpy.ndim(x) else out[0]


def _table(f):
    if isinstance(f, interpolate):
        return f
    data = getattr(f, '_data', None)           # a scipy InterpolatedUnivariateSpline keeps (x, y, ...) in _data
    if data is not None:
        return interpolate(data[0], data[1])
    raise TypeError('spline arguments must be table-backed (modi.nkk.interpolate or scipy InterpolatedUnivariateSpline)')


def N_inst(thetab, delT, l, mode='T'):
    ''' Return N_l spectrum given Delta_T [uK-arcmin] and beam-size [arcmin]'''
    arcm = (60 * 180 / numpy.pi) ** -1
    thetab *= arcm
    Tcmb = 2.786e6
    Nl = (delT * arcm / Tcmb) ** 2 * numpy.exp(l * (l + 1) * thetab ** 2. / (8 * numpy.log(2)))
    if mode == 'T':
        return Nl
    else:
        return 2. * Nl


def _double_integral(mode, Ls, l, tabs, nphi=200, device=None):
    ctx = _ffi.context(device)
    tabs = [_table(t) for t in tabs]
    l = _ffi.f64(l)
    phis = numpy.linspace(0, 2 * numpy.pi, nphi)
    w_phi = _ffi.f64(simps_weights_x(phis))            # simps(frac, phis, axis=-1), modi/nkk.py:88
    w_l = _ffi.f64(simps_weights_x(l))                 # simps(int1, l), :90
    cphi, sphi = _ffi.f64(numpy.cos(phis)), _ffi.f64(numpy.sin(phis))
    Ld = _ffi.f64(numpy.atleast_1d(Ls))
    tab_n = numpy.ascontiguousarray([t.y.size for t in tabs], dtype=numpy.int32)
    tab_x0 = _ffi.f64([t.x0 for t in tabs])
    tab_y = _ffi.f64(numpy.concatenate([t.y for t in tabs]))
    out = numpy.empty(Ld.size)
    _ffi.check(_ffi.lib().lbg_n0_modi(ctx.handle, MODE_BITS[mode], Ld.size, _ffi._dp(Ld), l.size, _ffi._dp(l), _ffi._dp(w_l),
                                      nphi, _ffi._dp(cphi), _ffi._dp(sphi), _ffi._dp(w_phi),
                                      tab_n.ctypes.data_as(_ffi.c_int_p), _ffi._dp(tab_x0), _ffi._dp(tab_y), _ffi._dp(out)))
    ms, launches = ctx.last_kernel_ms()
    last_stats.update(kernel_ms=ms, launches=launches, nodes=Ld.size * l.size * nphi)
    return out


def N_fisher(L, l, icl1, icl2, icln1, icln2, mode):
    '''Do the variance double integral for TT, EE and EB; L may be a scalar (as in the reference) or an array.'''
    if mode not in ('TT', 'EE', 'EB'):
        raise ValueError('N_fisher handles TT, EE and EB; use N_fisherTE for TE')
    out = _double_integral(mode, L, l, (icl1, icl2, icln1, icln2))
    return out if numpy.ndim(L) else out[0]


def N_fisherTE(L, l, iclte, iclnt, iclne, iclnte):
    '''Do the variance double integral for TE'''
    out = _double_integral('TE', L, l, (iclte, iclnt, iclne, iclnte))
    return out if numpy.ndim(L) else out[0]


class CMB:
    def __init__(self, clfile="RunPB02_cl.dat", clsfile="RunPB02_cl_lensed.dat"):
        self.clfile = clfile
        self.clsfile = clsfile
        self.read_class()
        self.lint = numpy.linspace(1., 3000., 10000)
        self.lintP = numpy.linspace(1., 5000., 10000)
        self.lat = numpy.logspace(2.0, 3.5, 15)
        self.kappa = self.cl[:, 5] * self.ll ** 2 / 4.
        self.kappas = self.cls[:, 5] * self.lls ** 2 / 4.

    def read_class(self):
        '''CLASS output for unlensed (clfile) and lensed (clsfile) spectra; columns l, TT, EE, TE, BB, phiphi, TPhi, Ephi.'''
        self.cols = ['l', 'TT', 'EE', 'TE', 'BB', 'phiphi', 'TPhi', 'Ephi']
        self.coldict = dict((name, i) for i, name in enumerate(self.cols))
        self.cl = numpy.loadtxt(self.clfile)
        self.cls = numpy.loadtxt(self.clsfile)
        self.ell = self.cl[:, 0]
        self.ells = self.cls[:, 0]
        self.ll = self.ell * (self.ell + 1)
        self.lls = self.ells * (self.ells + 1)
        self.tpi = 2 * numpy.pi
        self.cl[:, 1:] = self.cl[:, 1:] * (self.tpi / self.ll.reshape(-1, 1))          # strip l(l+1)/2pi
        self.cls[:, 1:] = self.cls[:, 1:] * (self.tpi / self.lls.reshape(-1, 1))

    def do_noise(self, mode1, mode2, thetab, delT):
        ''' Return Fisher noise for mode = mode1 + mode2 '''
        auto1 = mode1 + mode1
        auto2 = mode2 + mode2
        if mode1 + mode2 == "TE" or mode1 + mode2 == "ET":
            print("Wrong function")
            return None
        icl1 = interpolate(self.ell, self.cl[:, self.coldict[auto1]])
        icl2 = interpolate(self.ell, self.cl[:, self.coldict[auto2]])
        Nl1 = DetectorNoise(self.ells, thetab, delT, auto1)
        Nl2 = DetectorNoise(self.ells, thetab, delT, auto2)
        icln1 = interpolate(self.ells, self.cls[:, self.coldict[auto1]] + Nl1)
        icln2 = interpolate(self.ells, self.cls[:, self.coldict[auto2]] + Nl2)
        if mode1 + mode2 == "EB" or mode1 + mode2 == "EE":
            print('Go to higher lmax for %s' % (mode1 + mode2))
            noise = N_fisher(self.lat, self.lintP, icl1, icl2, icln1, icln2, mode1 + mode2)   # all L in one launch
        else:
            noise = N_fisher(self.lat, self.lint, icl1, icl2, icln1, icln2, mode1 + mode2)
        noise = (2 * numpy.pi) ** 2. / noise
        return Nl1, Nl2, noise * (self.lat * (self.lat + 1)) ** 2. / 4.

    def do_noiseTE(self, thetab, delT):
        '''Return Fisher noise for TE'''
        icl3 = interpolate(self.ell, self.cl[:, self.coldict["TE"]])
        Nl1 = DetectorNoise(self.ells, thetab, delT, "TT")
        Nl2 = DetectorNoise(self.ells, thetab, delT, "EE")
        icln1 = interpolate(self.ells, self.cls[:, self.coldict["TT"]] + Nl1)
        icln2 = interpolate(self.ells, self.cls[:, self.coldict["EE"]] + Nl2)
        icln3 = interpolate(self.ells, self.cls[:, self.coldict["TE"]])                        # NO noise added to TE
        noise = N_fisherTE(self.lat, self.lint, icl3, icln1, icln2, icln3)
        noise = (2 * numpy.pi) ** 2. / noise
        return Nl1, Nl2, noise * (self.lat * (self.lat + 1)) ** 2. / 4.
