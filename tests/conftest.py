import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref/libref_host.so (reference sources built for the host)")


@pytest.fixture(scope="session")
def rt():
    """The product package (directory name has hyphens, hence importlib)."""
    return importlib.import_module("ray-tracing-cuda_b200")
