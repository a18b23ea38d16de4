"""Import the UNMODIFIED reference modules from /root/reference (build container only).

TEST INFRASTRUCTURE.  Used by oracle/make_golden.py and by the `-m "not gpu"` pinning tests
(skipped when /root/reference is absent, i.e. on the GPU box).  Nothing in the reference tree is
copied or modified; this only arranges sys.path / stubs so its py2-era modules import on
python 3.12 / numpy 2 / scipy 1.18.
"""
import builtins
import functools
import os
import sys
import tempfile

REF = os.environ.get('LBGCMB_REFERENCE', '/root/reference')
_done = False


def available():
    return os.path.isfile(os.path.join(REF, 'cmb', 'Nkk.py'))


def setup():
    global _done
    if _done:
        return
    if not available():
        raise RuntimeError('reference tree not present at %s' % REF)
    import numpy as np
    import scipy.integrate
    from . import simps121
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path[:0] = [os.path.join(here, 'shim'), REF, os.path.join(REF, 'cmb'), os.path.join(REF, 'modi')]
    import matplotlib  # noqa: F401  (the shim registers matplotlib.pyplot and pylab)
    # python-2 / old-numpy isms used by the reference
    for name, typ in (('int', int), ('float', float)):
        if not hasattr(np, name):
            setattr(np, name, typ)
    if not hasattr(np, 'trapz'):
        np.trapz = np.trapezoid
    builtins.reduce = functools.reduce
    # scipy 1.2.1 simps -- NOT scipy.integrate.simpson
    scipy.integrate.simps = simps121.simps
    os.environ.setdefault('LBGCMB', tempfile.mkdtemp(prefix='lbgcmb_scratch_'))
    _done = True


def load(name):
    """import one reference module by its own top-level name (e.g. 'Nkk', 'Cgg', 'halofit')."""
    setup()
    import importlib
    return importlib.import_module(name)
