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
@pytest.mark.parametrize("exe", ["test_facade", "test_module", "test_group"])
def test_cpp_boundary_on_gpu(built, exe):
    """dlib's KAT / golden values through the facade; the module driven headless, bit-equal to the CPU glue."""
    r = subprocess.run([os.path.join(built, exe)], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "0 failed" in r.stdout


# ------------------------------------------------------------------------------------------------------------------
# Cycle-by-cycle comparison of three builds of the MODULE library over one scenario (tests/cpp/module_driver.cpp).
# Golden = the reference's OWN module source + the real dlib::mpc (oracle/_ref/libref_module_dlib.so, CPU), recorded
# in tests/golden/module_cycles.txt by:  tests/cpp/build/module_driver oracle/_ref/libref_module_dlib.so
GOLDEN_CYCLES = os.path.join(ROOT, "tests", "golden", "module_cycles.txt")
REF_MODULE_DLIB = os.path.join(ROOT, "oracle", "_ref", "libref_module_dlib.so")
REF_MODULE_BMPC = os.path.join(BUILD, "libref_module_bmpc.so")
OUR_MODULE = os.path.join(CPP, "build", "libtrajectory_point_controller.so")


def _drive(built, module_path):
    r = subprocess.run([os.path.join(built, "module_driver"), module_path], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    return r.stdout.splitlines()


def test_golden_cycles_come_from_the_reference_module(built):
    """Pins the golden file: the unmodified reference module + real dlib reproduces it (needs /root/reference)."""
    if not os.path.exists(REF_MODULE_DLIB):
        pytest.skip("oracle/_ref/libref_module_dlib.so not built (no /root/reference on this machine)")
    assert _drive(built, REF_MODULE_DLIB) == open(GOLDEN_CYCLES).read().splitlines()


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["rehosted_module", "reference_source_on_facade"])
def test_module_cycles_match_reference(built, which):
    """CAR / TRAJECTORY_POINT outputs of 48 cycles (tobiMPC incl. FOH look-ahead, crossing slow-down, lane-change
    indicators, IDLE, PID fallback) are identical — every float bit — to the reference module on the CPU."""
    path = OUR_MODULE if which == "rehosted_module" else REF_MODULE_BMPC
    if not os.path.exists(path):
        pytest.skip(f"{path} was not prebuilt (needs /root/reference at build time)")
    got, want = _drive(built, path), open(GOLDEN_CYCLES).read().splitlines()
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g == w
