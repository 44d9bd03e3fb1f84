import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds on the CPU")


@pytest.fixture(scope="session")
def ref_bpmodel():
    """The compiled UNMODIFIED reference extension (oracle/_ref), or skip."""
    from oracle import refenv
    mod = refenv.load_ref_bpmodel()
    if mod is None:
        pytest.skip("oracle/_ref reference extension not built")
    return mod
