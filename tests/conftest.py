import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def built():
    """Build (if stale) the CUDA library and the oracle once per session."""
    from rust_raytracer_b200.build import build_library, build_oracle
    build_library()
    build_oracle()
    return True


@pytest.fixture(scope="session")
def ctx(built):
    import rust_raytracer_b200 as rt
    return rt.Context(0)
