"""The C-ABI library loads and exports every symbol include/bmpc.h declares; error paths that need no GPU.
[-m "not gpu"]"""
import ctypes as C
import os
import re
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    hdr = open(os.path.join(ROOT, "include", "bmpc.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return sorted(set(re.findall(r"\b(bmpc_[a-z0-9_]+)\s*\(", hdr)))


def test_header_symbols_exported(engine_lib):
    L = engine_lib.lib()
    names = _declared_symbols()
    assert len(names) >= 20
    for n in names:
        assert hasattr(L, n), f"libbmpc.so does not export {n}"
    # ... and the binding knows each of them with a signature
    assert set(names) == set(engine_lib.SIGNATURES)


def test_exports_are_plain_c(engine_lib):
    out = subprocess.run(["nm", "-D", "--defined-only", engine_lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    for n in _declared_symbols():
        assert n in exported


def test_shape_table(engine_lib):
    for s in [(2, 2, 4), (2, 1, 30), (4, 1, 10), (4, 1, 30), (4, 1, 60), (4, 1, 100)]:
        assert engine_lib.shape_supported(*s)
    assert not engine_lib.shape_supported(5, 1, 30)
    assert b"sm_100a" in engine_lib.lib().bmpc_version()


def test_create_rejects_bad_arguments(engine_lib):
    L = engine_lib.lib()
    h = C.c_void_p()
    assert L.bmpc_create(C.byref(h), 5, 1, 30, 16, 64, 0, -1) == -2          # shape not instantiated
    assert b"not instantiated" in L.bmpc_last_error()
    assert L.bmpc_create(C.byref(h), 4, 1, 30, 0, 64, 0, -1) == -1           # capacity
    assert L.bmpc_create(C.byref(h), 4, 1, 30, 16, 16, 0, -1) == -2          # precision
    assert L.bmpc_create(None, 4, 1, 30, 16, 64, 0, -1) == -1


def test_no_cpu_fallback_without_device(engine_lib):
    """On a machine without a GPU the product path must fail loudly, not compute on the CPU."""
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if has_gpu:
        pytest.skip("a GPU is present")
    with pytest.raises(engine_lib.BmpcError) as e:
        engine_lib.BatchedMPC(4, 1, 30, 16)
    assert e.value.code == -3 and "no CPU fallback" in str(e.value)


def test_group_create_rejects_bad_arguments(engine_lib):
    """Device groups: argument errors and the loud no-device failure need no GPU."""
    L = engine_lib.lib()
    g = C.c_void_p()
    assert L.bmpc_group_create(None, 4, 1, 30, 64, 64, 0, None, 0, 0) == -1
    assert L.bmpc_group_create(C.byref(g), 4, 1, 30, 64, 64, 0, None, 0, 7) == -1          # unknown assignment
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except Exception:
        has_gpu = False
    if not has_gpu:
        assert L.bmpc_group_create(C.byref(g), 4, 1, 30, 64, 64, 0, None, 0, 0) == -3
        assert b"no CPU fallback" in L.bmpc_last_error()
    assert L.bmpc_group_size(None) == 0 and L.bmpc_group_member(None, 0) is None


def test_product_never_imports_oracle():
    """Nothing under trajectory_controller_b200/ may reference oracle/ (SURVEY.md §8(c) rule)."""
    pkg = os.path.join(ROOT, "trajectory_controller_b200")
    for dp, _, files in os.walk(pkg):
        if os.sep + "build" in dp:
            continue
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                src = open(os.path.join(dp, f), errors="ignore").read()
                assert "liboracle" not in src and "_oracle" not in src and "libdlibref" not in src, os.path.join(dp, f)
