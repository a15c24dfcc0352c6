"""pytest configuration: the `gpu` marker, and handles to the product package, the CPU oracle
(oracle/libpm_oracle.so) and — where it was built — the reference-derived checker (oracle/_ref)."""
import ctypes
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

import __graft_entry__ as graft  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def pkg():
    return graft.load_package()


@pytest.fixture(scope="session")
def orc():
    from helpers import load_oracle

    return load_oracle()


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref/libcukd_ref.so: the reference's own cukd/OWL code behind a C shim (may be absent)."""
    path = os.path.join(ROOT, "oracle", "_ref", "libcukd_ref.so")
    if not os.path.exists(path):
        pytest.skip("oracle/_ref not built (needs /root/reference at build time)")
    L = ctypes.CDLL(path)
    L.ref_lcg.restype = ctypes.c_uint32
    return L


@pytest.fixture(scope="session")
def ctx(pkg):
    import torch

    assert torch.cuda.is_available(), "gpu tests need a CUDA device"
    c = pkg.Context(0)
    yield c
    c.close()
