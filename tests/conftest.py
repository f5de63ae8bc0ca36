import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    """The CPU oracle library (test infrastructure)."""
    from oracle import loader

    return loader.load()


@pytest.fixture(scope="session")
def product():
    """The CUDA product library; GPU tests only."""
    from fimo_b200 import library

    return library.load()
