"""The host C++ side above the C ABI: dlib::mpc-compatible facade and the re-hosted LMS module
(trajectory_controller_b200/cpp).  The binaries are built by cpp/Makefile (also by __graft_entry__.build())."""
import ctypes
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CPP = os.path.join(ROOT, "trajectory_controller_b200", "cpp")
TESTS_CPP = os.path.join(ROOT, "tests", "cpp")
BUILD = os.path.join(TESTS_CPP, "build")


@pytest.fixture(scope="module")
def built():
    from _oracle import Oracle
    Oracle()                                                    # makes sure oracle/liboracle.so exists (test_module links it)
    subprocess.run(["make", "-s", "-C", CPP], check=True)
    subprocess.run(["make", "-s", "-C", TESTS_CPP], check=True)
    return BUILD


def test_module_library_exports_the_lms_factory(built):
    """The MODULE library exposes the C factory the LMS loader dlsym()s (reference src/interface.cpp:3)."""
    lib = ctypes.CDLL(os.path.join(CPP, "build", "libtrajectory_point_controller.so"))
    assert hasattr(lib, "getInstance") and hasattr(lib, "getLmsVersion")
    lib.getLmsVersion.restype = ctypes.c_uint32
    assert lib.getLmsVersion() > 0


def test_no_silent_cpu_fallback(built):
    """Without a CUDA device the facade throws and the module refuses to initialise — loudly."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the loud-failure path is exercised on CPU-only machines")
    for exe, needle in (("test_facade", "no CUDA device"), ("test_module", "initialize() failed")):
        r = subprocess.run([os.path.join(built, exe)], capture_output=True, text=True, timeout=120)
        assert r.returncode == 2 and needle in r.stderr, (exe, r.returncode, r.stderr[-300:])


@pytest.mark.gpu
@pytest.mark.parametrize("exe", ["test_facade", "test_module"])
def test_cpp_boundary_on_gpu(built, exe):
    """dlib's KAT / golden values through the facade; the module driven headless, bit-equal to the CPU glue."""
    r = subprocess.run([os.path.join(built, exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "0 failed" in r.stdout
