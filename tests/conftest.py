import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle (oracle/pf_oracle.c), built on first use."""
    from oracle import pf_oracle
    pf_oracle.lib()
    return pf_oracle


@pytest.fixture(scope="session")
def pf():
    """The product package; builds libpillowfight_b200.so if it is stale."""
    from __graft_entry__ import load_build_module
    load_build_module().build()
    import libpillowfight_b200
    return libpillowfight_b200
