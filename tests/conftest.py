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


_probe = {}


@pytest.hookimpl(trylast=True)
def pytest_runtest_teardown(item, nextitem):
    """B200RT_PROBE_ERRORS=1 (diagnostic): after every GPU test, one 1-ray b200rt_intersect_batch on a private context.  That entry
    point checks cudaGetLastError() right after its launch, so a non-sticky CUDA error an earlier call left behind in this
    thread surfaces HERE, next to the test that caused it, instead of inside whichever API call polls the error state next."""
    if not os.environ.get("B200RT_PROBE_ERRORS") or "gpu" not in item.keywords:
        return
    import numpy as np
    import rust_raytracer_b200 as rt
    if not _probe:
        from scenes_util import scene_builder
        c = rt.Context(0)
        _probe.update(ctx=c, scene=rt.Scene(c, scene_builder("cornell_ref")))
    o = np.array([[278, 278, -800, 0]], np.float32)
    d = np.array([[0, 0, 1, np.inf]], np.float32)
    try:
        _probe["scene"].intersect_batch(o, d)
    except rt.B200RTError as e:
        sys.stderr.write("\n[probe] stale CUDA error after %s: %s\n" % (item.nodeid, e))
