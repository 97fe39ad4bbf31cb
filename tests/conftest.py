import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def oracle():
    import oracle_lib
    return oracle_lib.Lib("oracle")


@pytest.fixture(scope="session")
def reflib():
    import oracle_lib
    if not oracle_lib.have_ref():
        pytest.skip("compiled reference (oracle/_ref/liboh_ref.so) not available and /root/reference absent")
    return oracle_lib.Lib("ref")
