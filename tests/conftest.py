import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a CUDA device (run on the B200 box)')


@pytest.fixture(scope='session')
def oracle_lib():
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope='session')
def wtlib():
    from wavelettransforms_b200 import build, _lib
    build.build_library()
    return _lib
