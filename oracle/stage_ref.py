"""Stage the reference's OWN Python implementation of the hot path for the GPU box (TEST / BASELINE INFRASTRUCTURE).

/root/reference does not travel to the GPU box; byte-compiled modules do (oracle/_ref/ is git-ignored but not gpurun-ignored,
like oracle/_ref/libeh_ref.so).  This recipe compiles the reference's path modules FROM WHERE THEY LIE to sourceless bytecode and
packs it into ONE archive, oracle/_ref/refpy.zip (same relative layout inside: cmb/Nkk.pyc, Cgg.pyc, ...; python imports
sourceless .pyc members of a zip on sys.path).  No source is copied into the repository and nothing is modified;
oracle/refload.py imports the staged modules when /root/reference is absent, so `bench.py --impl reference` times the UNMODIFIED
reference (cmb/Nkk.py, Cgg.py, Ckg.py, cmb/lensing.py) on the box's host cores.

usage: python oracle/stage_ref.py [reference root]      (run by `make -C oracle`, i.e. by __graft_entry__.build())"""
import os
import py_compile
import sys

MODULES = ['Cgg.py', 'Ckg.py', 'cmb/__init__.py', 'cmb/Nkk.py', 'cmb/lensing.py', 'cmb/DetectorNoise.py', 'pmh.py', 'pmm.py',
           'prep_camb.py', 'utils.py', 'cosmo.py', 'growth_rate.py', 'sliced_pz.py', 'pz2nbar.py', 'params.py', 'Gaussian_pz.py',
           'prep_Llls.py']


def main(ref='/root/reference'):
    import tempfile
    import warnings
    import zipfile
    here = os.path.dirname(os.path.abspath(__file__))
    dest = os.path.join(here, '_ref', 'refpy.zip')
    if not os.path.isfile(os.path.join(ref, 'cmb', 'Nkk.py')):
        print('stage_ref: %s absent, nothing staged' % ref)
        return 0
    warnings.simplefilter('ignore', SyntaxWarning)
    os.makedirs(os.path.dirname(dest), exist_ok=True)
    with tempfile.TemporaryDirectory() as tmp, zipfile.ZipFile(dest, 'w', zipfile.ZIP_DEFLATED) as z:
        for rel in MODULES:
            out = os.path.join(tmp, rel[:-3] + '.pyc')
            os.makedirs(os.path.dirname(out), exist_ok=True)
            py_compile.compile(os.path.join(ref, rel), cfile=out, dfile='reference/' + rel, doraise=True)
            z.write(out, rel[:-3] + '.pyc')
        z.writestr('STAGED_FROM', '%s (python %d.%d bytecode; sources not copied)\n' % (ref, sys.version_info[0], sys.version_info[1]))
    print('staged %d reference modules in oracle/_ref/refpy.zip' % len(MODULES))
    return 0


if __name__ == '__main__':
    sys.exit(main(*sys.argv[1:2]))
