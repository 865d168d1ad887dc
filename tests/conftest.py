import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box with -m gpu)')


@pytest.fixture(scope='session')
def built():
    """Both native libraries: the product (.so for sm_100a) and the oracle (test infrastructure)."""
    from urgym_b200 import build
    import oracle_lib
    build.build()
    oracle_lib.build()
    return True
