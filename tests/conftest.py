import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built_lib():
    """Build (if stale) and load the C-ABI library; CPU-safe (no compute calls)."""
    import cloud_server_b200 as pkg
    from cloud_server_b200 import _lib
    pkg.build_library()
    return _lib.lib()


@pytest.fixture(scope="session")
def ctx(built_lib):
    from cloud_server_b200.context import Context
    return Context(0)


def golden(name):
    import numpy as np
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
