import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: a BASELINE-size case (tens of seconds: the reference CPU factorization beside ours)")


@pytest.fixture(scope="session", autouse=True)
def _built():
    """Make sure the product library and the oracle exist (compiles them if a toolchain is here)."""
    import __graft_entry__ as g
    g.build()
