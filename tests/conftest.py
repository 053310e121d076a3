import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG_NAME = "surface-irradiance-estimation-from-neural-radiance-fields_b200"
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def hn():
    lib = os.path.join(ROOT, PKG_NAME, "libhn_b200.so")
    if not os.path.exists(lib):
        import __graft_entry__ as g
        g.build()
    return importlib.import_module(PKG_NAME)


@pytest.fixture(scope="session")
def orc():
    import orc as _orc
    return _orc


@pytest.fixture(scope="session")
def root():
    return ROOT
