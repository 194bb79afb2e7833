import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")
    config.addinivalue_line("markers", "slow: takes more than a few seconds of CPU")


@pytest.fixture(scope="session")
def port_oracle():
    from oracle.oracle_api import Oracle, build, PORT_LIB
    if not os.path.exists(PORT_LIB):
        build(ref=os.path.exists("/root/reference/raytracing.c"))
    return Oracle("port")


@pytest.fixture(scope="session")
def ref_oracle():
    from oracle.oracle_api import Oracle, have_reference_lib
    if not have_reference_lib():
        pytest.skip("oracle/_ref/libref_oracle.so not built (needs /root/reference)")
    return Oracle("reference")
