import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def orc():
    """The CPU oracle (test infrastructure)."""
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def aslam():
    """The product package; the CUDA library must already be built (no fallback)."""
    import awesomeslam_b200
    from awesomeslam_b200 import build
    if not os.path.exists(awesomeslam_b200.capi.LIB_PATH):
        build.build_library()
    awesomeslam_b200.capi.lib()
    return awesomeslam_b200


@pytest.fixture(scope="session")
def aslam_host():
    """The package for HOST-side input generation only (maps, replay generator: lib/libaslam_host.so, no CUDA)."""
    import awesomeslam_b200
    from awesomeslam_b200 import build
    build.build_host_lib()
    return awesomeslam_b200
