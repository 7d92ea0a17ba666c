import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def table():
    from oracle_lib import load_table
    return load_table()


@pytest.fixture(scope="session")
def ref(table):
    from oracle_lib import RefOracle, REF_SO, build_oracles
    if not os.path.exists(REF_SO):
        build_oracles()
    if not os.path.exists(REF_SO):
        pytest.skip("oracle/_ref/libref_oracle.so not built and /root/reference absent")
    return RefOracle(table)


@pytest.fixture(scope="session")
def port(table):
    from oracle_lib import PortOracle
    return PortOracle(table)
