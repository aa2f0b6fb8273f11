import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def device():
    """One rp_handle on cuda:0 through the C ABI; fails loudly if the CUDA library cannot be used."""
    import rampack_b200 as rp
    dev = rp.Device(device=0, seed=1234)
    yield dev
    dev.close()
