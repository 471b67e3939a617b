"""pytest configuration: markers, import path, shared fixtures.

`-m "not gpu"`  : oracle vs golden vectors / real dlib, host logic, C-ABI symbol export.  Runs anywhere.
`-m gpu`        : parity of the CUDA path against the oracle, called through the C-ABI.  Needs a B200.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "tests")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "ref: needs oracle/_ref (the real dlib::mpc built from /root/reference)")


@pytest.fixture(scope="session")
def oracle():
    from _oracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def dlibref():
    from _oracle import DlibRef
    if not DlibRef.available():
        pytest.skip("oracle/_ref/libdlibref.so not built (no /root/reference on this machine)")
    return DlibRef()


@pytest.fixture(scope="session")
def golden():
    import numpy as np
    path = os.path.join(ROOT, "tests", "golden", "mpc_golden.npz")
    return np.load(path)


@pytest.fixture(scope="session")
def engine_lib():
    """The product C-ABI library (trajectory_controller_b200/libbmpc.so) through its ctypes binding."""
    from trajectory_controller_b200 import engine
    return engine
