import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, 'tests', 'golden')


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')
    config.addinivalue_line('markers', 'reference: needs /root/reference (build container only)')


@pytest.fixture(scope='session')
def golden_dir():
    return GOLDEN


@pytest.fixture(scope='session')
def ctx():
    """GPU context; fails loudly (no skip, no fallback) when the library or device is missing."""
    from lbgcmb_b200 import _ffi
    return _ffi.context(0)


@pytest.fixture(scope='session')
def cl_tables():
    from lbgcmb_b200.synthetic import cmb_cl_tables
    return cmb_cl_tables()
