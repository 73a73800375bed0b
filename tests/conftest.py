import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line('markers', 'gpu: needs a B200 (run with -m gpu on the GPU box)')


@pytest.fixture(scope='session')
def native_lib():
    """Path of the built shared library (built on demand; nvcc cross-compiles)."""
    from degenprimer_b200.build import build_native
    return build_native()


@pytest.fixture(scope='session')
def ctx(native_lib):
    """Native context on cuda:0 with the NN tables loaded (GPU tests only)."""
    from degenprimer_b200 import TD_Functions as tdf
    return tdf._ctx()
