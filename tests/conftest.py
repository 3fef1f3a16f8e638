import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def native():
    """The in-tree native library; built on demand (nvcc cross-compiles without a GPU)."""
    import helfem_b200
    if not os.path.exists(helfem_b200.LIB_PATH):
        from helfem_b200 import build
        build.build()
    return helfem_b200.lib()


@pytest.fixture(scope="session")
def oracle_lib():
    import oracle
    return oracle.lib()
