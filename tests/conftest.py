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


@pytest.fixture(autouse=True)
def _ieee_sphere_math(monkeypatch):
    """Sphere::Hit squares and roots through pow() (sphere.cu:13-21).  The product calls CUDA's pow like the reference's
    CUDA build (not correctly rounded: 1 ulp off x*x / sqrt on 6 % / 14 % of operands, bench.py pow_kat_device); the
    CPU oracle is pinned to the reference's HOST build, i.e. to correctly rounded values.  Tests that compare the CUDA
    path with the oracle bit for bit therefore commit their scenes with RT_SPHERE_POW=ieee (read at rt_scene_commit);
    tests/test_device_reference.py checks the default against the reference's own CUDA code."""
    monkeypatch.setenv("RT_SPHERE_POW", "ieee")
