import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: longer CPU test")


@pytest.fixture(scope="session")
def oracle():
    from tests import helpers
    return helpers.oracle_lib()


@pytest.fixture(scope="session")
def ref():
    from tests import helpers
    lib = helpers.ref_lib()
    if lib is None:
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return lib
