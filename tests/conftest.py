import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def built():
    """Native pieces built in-tree (nvcc cross-compiles without a GPU)."""
    import __graft_entry__ as ge
    ge.build()
    return ge


@pytest.fixture(scope="session")
def pkg(built):
    return built.load_package()


@pytest.fixture(scope="session")
def synth(built, pkg):
    return importlib.import_module(built.PKG_NAME + ".synth")


@pytest.fixture(scope="session")
def pipeline(built, pkg):
    return importlib.import_module(built.PKG_NAME + ".pipeline")


@pytest.fixture(scope="session")
def po(built):
    from oracle import pm_oracle
    pm_oracle.lib()
    return pm_oracle


@pytest.fixture()
def ctx(pkg):
    c = pkg.Context(0)
    yield c
    c.close()
