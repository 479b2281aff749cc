import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def backend():
    """The CUDA backend module; fails loudly (no fallback) when the library is missing."""
    import splatt3r_slam_b200.mast3r_slam_backends as be
    return be


@pytest.fixture(scope="session")
def ref_backend():
    """The reference CUDA build (oracle/_ref), or None when it did not travel to this box."""
    import importlib.util
    so = os.path.join(ROOT, "oracle", "_ref", "mast3r_slam_backends_ref.so")
    if not os.path.exists(so):
        return None
    try:
        spec = importlib.util.spec_from_file_location("mast3r_slam_backends_ref", so)
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
        return mod
    except Exception:
        return None
