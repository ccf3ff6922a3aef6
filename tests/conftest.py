import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
if HERE not in sys.path:
    sys.path.insert(0, HERE)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle

    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def pkg():
    """The ctypes harness over liblance_b200.so (the C ABI of include/lance_b200.h)."""
    from __graft_entry__ import load_package

    p = load_package()
    if not os.path.exists(p.LIB_PATH):
        p.build()
    return p


@pytest.fixture(scope="session")
def host(pkg):
    """ctypes harness over the C++ host mirror (csrc/host/, include/lance_b200_host.h)."""
    import importlib

    return importlib.import_module("duckdb_lancedb_b200.host")


@pytest.fixture(scope="session")
def gpu(pkg):
    """A CUDA device must be present for -m gpu tests: fail loudly, never skip to a fallback."""
    n = pkg.device_count()
    assert n >= 1, "no CUDA device: the gpu-marked tests must run on the B200 box"
    pkg.set_default_device(0)
    return pkg
