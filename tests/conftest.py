import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def _has_gpu():
    try:
        import delaunay_linterp_b200 as dl
        return dl.lib().dli_device_count() > 0
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu():
    """GPU tests must fail loudly, not skip, when selected with -m gpu on a box without the library."""
    import delaunay_linterp_b200 as dl
    n = dl.lib().dli_device_count()
    assert n > 0, "no CUDA device visible: the product has no CPU fallback"
    return n
