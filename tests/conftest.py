import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def ffemb():
    """The product package (its directory name has a hyphen)."""
    return importlib.import_module("dlrm-flexflow_b200")


@pytest.fixture(scope="session")
def oracle():
    """The parity checker (test infrastructure)."""
    from oracle import oracle as O
    O.lib()
    return O


def golden_path(name):
    return os.path.join(ROOT, "tests", "golden", name)
