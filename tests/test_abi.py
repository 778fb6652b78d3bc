"""CPU-only checks of the drop-in boundary: the library builds for sm_100a, loads, and exports every symbol
include/mazetd_b200.h declares. No compute call is made here (no GPU in this container)."""
import ctypes
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from mazetd_b200 import build

    path = build.build()
    return ctypes.CDLL(path)


def header_functions():
    src = open(os.path.join(ROOT, "include", "mazetd_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(mtd_[a-z0-9_]+)\s*\(", src)))


def test_header_declares_the_expected_surface():
    names = header_functions()
    from mazetd_b200 import capi

    assert sorted(capi.EXPORTS) == names


def test_library_exports_every_declared_symbol(lib):
    for name in header_functions():
        assert hasattr(lib, name), f"{name} declared in include/mazetd_b200.h but not exported"


def test_no_gpu_means_loud_failure(lib):
    """mtd_create must fail (not fall back to a CPU path) when there is no CUDA device."""
    import torch

    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = ctypes.c_void_p()
    lib.mtd_create.argtypes = [ctypes.POINTER(ctypes.c_void_p), ctypes.c_int, ctypes.c_void_p]
    rc = lib.mtd_create(ctypes.byref(h), 0, None)
    assert rc == -2 and not h.value  # MTD_ERR_CUDA
    from mazetd_b200 import capi

    with pytest.raises(capi.MtdError):
        capi.Context(0)


def test_product_does_not_import_the_oracle():
    """the product path must never route through oracle/ (SURVEY.md 8c / task rule 3)"""
    pkg = os.path.join(ROOT, "mazetd_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h", ".hpp")):
                text = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "oracle_lib" not in text and "libmazetd_oracle" not in text and "libmazetd_ref" not in text, f
                assert not re.search(r"#include\s+[\"<].*oracle/", text), f


def test_sass_is_sm100a(lib):
    from mazetd_b200 import build
    import subprocess

    out = subprocess.run(["/usr/local/cuda/bin/cuobjdump", "-lelf", build.LIB], capture_output=True, text=True).stdout
    assert "sm_100a" in out
