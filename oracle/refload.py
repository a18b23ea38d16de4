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
# the reference's path modules byte-compiled by oracle/stage_ref.py (what travels to the GPU box instead of /root/reference)
STAGED = os.path.join(os.path.dirname(os.path.abspath(__file__)), '_ref', 'refpy.zip')
_done = False


def available():
    """The full reference tree (build container)."""
    return os.path.isfile(os.path.join(REF, 'cmb', 'Nkk.py'))


def staged_available():
    """The byte-compiled path modules (cmb/Nkk, Cgg, Ckg, cmb/lensing and their imports): enough for the CPU baseline arm."""
    return os.path.isfile(STAGED)


def root():
    return REF if available() else (STAGED if staged_available() else None)


def setup():
    global _done
    if _done:
        return
    if root() is None:
        raise RuntimeError('reference tree not present at %s (and nothing staged under %s)' % (REF, STAGED))
    import numpy as np
    import scipy.integrate
    from . import simps121, romberg121
    here = os.path.dirname(os.path.abspath(__file__))
    base = root()
    sys.path[:0] = [os.path.join(here, 'shim'), base, os.path.join(base, 'cmb'), os.path.join(base, 'modi')]
    import matplotlib  # noqa: F401  (the shim registers matplotlib.pyplot and pylab)
    # python-2 / old-numpy isms used by the reference
    for name, typ in (('int', int), ('float', float)):
        if not hasattr(np, name):
            setattr(np, name, typ)
    if not hasattr(np, 'trapz'):
        np.trapz = np.trapezoid
    if not hasattr(np, 'NaN'):
        np.NaN = np.nan                      # dropouts/bz/py/get_bz.py:43-44
    builtins.reduce = functools.reduce
    # scipy 1.2.1 simps -- NOT scipy.integrate.simpson
    scipy.integrate.simps = simps121.simps
    # scipy 1.2.1 romberg (modi/cosmology.py:6,107); removed from scipy >= 1.15
    scipy.integrate.romberg = romberg121.romberg
    os.environ.setdefault('LBGCMB', tempfile.mkdtemp(prefix='lbgcmb_scratch_'))
    _done = True


def load(name):
    """import one reference module by its own top-level name (e.g. 'Nkk', 'Cgg', 'halofit')."""
    setup()
    import importlib
    return importlib.import_module(name)


def load_modi():
    """(cosmology, lensing): the reference's modi/cosmology.py and modi/lensing.py, loaded from where they lie.

    They import their siblings under `modi_*` names that exist nowhere in the tree (modi/cosmology.py:10-11,
    modi/lensing.py:6-9) and modi/lensing.py carries two python-2 print statements (:83, :294): the aliases are registered
    here and the two statements are parenthesised IN MEMORY before compiling -- nothing is copied or written."""
    setup()
    import re
    import types
    if 'modi_lensing_ref' in sys.modules:
        return sys.modules['modi_cosmology'], sys.modules['modi_lensing_ref']
    sys.modules['modi_halofit'] = load('halofit')
    for stub in ('modi_tools', 'modi_pk_pregen'):            # circular helper imports; unused by Cosmology / LensingX
        sys.modules.setdefault(stub, types.ModuleType(stub))
    def exec_file(modname, path, fix=None):
        src = open(path).read()
        if fix:
            src = fix(src)
        mod = types.ModuleType(modname)
        mod.__file__ = path
        sys.modules[modname] = mod
        exec(compile(src, path, 'exec'), mod.__dict__)
        return mod
    cosmology = exec_file('modi_cosmology', os.path.join(REF, 'modi', 'cosmology.py'))
    py3 = lambda src: re.sub(r'^(\s*)print (.+)$', lambda m: '%sprint(%s)' % (m.group(1), m.group(2)), src, flags=re.M)
    lensing = exec_file('modi_lensing_ref', os.path.join(REF, 'modi', 'lensing.py'), fix=py3)
    return cosmology, lensing
