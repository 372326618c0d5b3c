import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_lib():
    from oracle import native
    native.build(with_ref=True)
    return native.load("best")


@pytest.fixture(scope="session")
def port_lib():
    from oracle import native
    native.build(with_ref=False)
    return native.load("port")
