import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def built():
    """Native libraries: built once per session if missing (nvcc / g++ cross-compile on CPU)."""
    from pooraytracer_b200 import _capi
    import oracle_lib as ol
    if not os.path.exists(_capi.LIB_PATH) or not ol.available("orc"):
        import __graft_entry__ as g
        g.build()
    return True


@pytest.fixture(scope="session")
def gpu_ctx(built):
    from pooraytracer_b200 import Context
    return Context(0)
