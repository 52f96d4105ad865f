import os
import sys
import warnings

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

warnings.filterwarnings("ignore")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "reference: needs the live reference under /root/reference")


@pytest.fixture(scope="session")
def ref():
    from oracle import ref_shim
    if not ref_shim.available():
        pytest.skip("live reference not present (GPU box): golden fixtures cover this")
    return ref_shim.load()
