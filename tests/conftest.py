import os
import sys

import pytest

os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")  # see umx.cpp_b200/__init__.py

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: minutes of CPU oracle time")


@pytest.fixture(scope="session")
def umx():
    """The product package (umx.cpp_b200), loaded by path."""
    import __graft_entry__ as g

    return g.load_package()


@pytest.fixture(scope="session")
def have_gpu(umx):
    return umx.lib().umx_cuda_device_count() > 0
