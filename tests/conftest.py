import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "reference: needs the reference tree at /root/reference (build container)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import salp_oracle
    salp_oracle.build()
    return salp_oracle


@pytest.fixture(scope="session")
def salp_lib():
    from underwater_swimmer_rl_b200 import build as b
    b.build()
    from underwater_swimmer_rl_b200 import _capi
    return _capi.load()
