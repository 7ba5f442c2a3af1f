import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    import __graft_entry__ as ge
    if not os.path.exists(ge.LIB):
        ge.build_product()
    return ge.load_package()


@pytest.fixture(scope="session")
def ob():
    import __graft_entry__ as ge
    ge.build_oracle()
    import oracle_bind
    return oracle_bind


@pytest.fixture(scope="session")
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()
