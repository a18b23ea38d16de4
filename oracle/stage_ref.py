"""Stage the reference's OWN Python implementation of the hot path for the GPU box (TEST / BASELINE INFRASTRUCTURE).

/root/reference does not travel to the GPU box; byte-compiled modules do (oracle/_ref/ is git-ignored but not gpurun-ignored,
like oracle/_ref/libeh_ref.so).  This recipe compiles the reference's path modules FROM WHERE THEY LIE to sourceless .pyc files
under oracle/_ref/refpy/ (same relative layout: cmb/Nkk.pyc, Cgg.pyc, ...).  No source is copied into the repository and nothing
is modified; oracle/refload.py imports the staged modules when /root/reference is absent, so `bench.py --impl reference` times
the UNMODIFIED reference (cmb/Nkk.py, Cgg.py, Ckg.py, cmb/lensing.py) on the box's host cores.

usage: python oracle/stage_ref.py [reference root]      (run by `make -C oracle`, i.e. by __graft_entry__.build())"""
import os
import py_compile
import sys

MODULES = ['Cgg.py', 'Ckg.py', 'cmb/__init__.py', 'cmb/Nkk.py', 'cmb/lensing.py', 'cmb/DetectorNoise.py', 'pmh.py', 'pmm.py',
           'prep_camb.py', 'utils.py', 'cosmo.py', 'growth_rate.py', 'sliced_pz.py', 'pz2nbar.py', 'params.py', 'Gaussian_pz.py',
           'prep_Llls.py']


def main(ref='/root/reference'):
    here = os.path.dirname(os.path.abspath(__file__))
    dest = os.path.join(here, '_ref', 'refpy')
    if not os.path.isfile(os.path.join(ref, 'cmb', 'Nkk.py')):
        print('stage_ref: %s absent, nothing staged' % ref)
        return 0
    import warnings
    warnings.simplefilter('ignore', SyntaxWarning)
    for rel in MODULES:
        out = os.path.join(dest, rel[:-3] + '.pyc')
        os.makedirs(os.path.dirname(out), exist_ok=True)
        py_compile.compile(os.path.join(ref, rel), cfile=out, dfile='reference/' + rel, doraise=True)
    with open(os.path.join(dest, 'STAGED_FROM'), 'w') as f:
        f.write('%s (python %d.%d bytecode; sources not copied)\n' % (ref, sys.version_info[0], sys.version_info[1]))
    print('staged %d reference modules under oracle/_ref/refpy' % len(MODULES))
    return 0


if __name__ == '__main__':
    sys.exit(main(*sys.argv[1:2]))
