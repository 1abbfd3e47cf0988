import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def built():
    """Builds the product library (nvcc cross-compiles without a GPU), the oracle and the generator."""
    import __graft_entry__ as g
    g.build()
    return True


@pytest.fixture(scope="session")
def oracle_lib(built):
    from oracle import pyorc
    return pyorc


@pytest.fixture(scope="session")
def synth(built):
    from tools.synth import synth as s
    return s


def _has_gpu():
    try:
        from lidarslam_b200 import capi
        return capi.load_library().lsb_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu(built):
    if not _has_gpu():
        pytest.fail("no CUDA device: -m gpu tests must run on a GPU box (there is no CPU fallback)")
    from lidarslam_b200.slam import Slam
    return Slam
