import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: test needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: takes tens of seconds (full-size parity checks)")


@pytest.fixture(scope="session")
def ref():
    """The reference's own collision code (oracle/_ref); tests that need it are skipped when it was not built."""
    from oracle import pyoracle

    if not pyoracle.have_ref():
        pytest.skip("oracle/_ref/libchrono_ref.so not built")
    return pyoracle.Ref()
