import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def load_product():
    """the product package; its directory name has dashes, so it is imported by string"""
    return importlib.import_module("vote-saver-protocol_b200")


@pytest.fixture(scope="session")
def vsp():
    return load_product()


@pytest.fixture(scope="session")
def ctx(vsp):
    # GPU tests must run on the CUDA path or fail: no skip on a missing device / library
    c = vsp.Context(0)
    yield c
    c.close()
