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
def oracle():
    """the strongest oracle present: the reference's own code compiled in place, else the C++ port"""
    import oracle_lib as O
    for kind in ("reference", "port"):
        if O.available(kind):
            return O.Oracle(kind)
    pytest.fail("no oracle library built: run `make -C oracle`")
