import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def snb():
    from synerfgine_b200 import _lib
    return _lib.load()


@pytest.fixture(scope="session")
def orc():
    import oracle_lib
    return oracle_lib.load()


@pytest.fixture(autouse=True)
def _oracle_host_scales(request):
    """GPU parity tests hand the device-evaluated grid scales to the oracle (common.use_device_grid_scales); every test starts and ends
    on the oracle's own host formula."""
    yield
    if request.node.get_closest_marker("gpu") is not None:
        import oracle_lib
        oracle_lib.load().orc_set_grid_scales(None, 0)
