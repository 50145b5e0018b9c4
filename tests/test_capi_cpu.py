"""CPU suite, part 2: the C-ABI library loads and exports every symbol include/gravidy_b200.h declares,
and refuses to run without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from gravidy_b200 import build as B
from gravidy_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "gravidy_b200.h")


@pytest.fixture(scope="module")
def lib():
    B.build()
    return capi.load()


def _declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(gdy_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_all_exported(lib):
    syms = _declared_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/gravidy_b200.h but not exported"
    # and the Python prototypes cover exactly the header
    assert sorted(capi.SIGNATURES) == syms


def test_abi_version(lib):
    assert lib.gdy_abi_version() == 2


def test_no_cpu_fallback(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    h = capi._ctx()
    rc = lib.gdy_create(C.byref(h), 16, 1e-8, C.c_float(0.01), 0)
    assert rc == capi.GDY_ERR_NO_DEVICE
    assert not h.value
    assert b"no CPU path" in lib.gdy_last_error(None)


def test_bad_arguments_rejected_before_device_work(lib):
    h = capi._ctx()
    assert lib.gdy_create(C.byref(h), 0, 1e-8, C.c_float(0.01), 0) == capi.GDY_ERR_INVALID
    assert lib.gdy_create_sharded(C.byref(h), 16, 1e-8, C.c_float(0.01), 0, 2, 2, None) == capi.GDY_ERR_INVALID
    assert lib.gdy_destroy(None) == 0
    assert lib.gdy_launch_count(None) == 0
    assert lib.gdy_sync(None) == capi.GDY_ERR_INVALID


def test_product_never_imports_oracle():
    """The product package must not import, link or dlopen anything under oracle/."""
    pkg = os.path.join(ROOT, "gravidy_b200")
    bad = re.compile(r"(^\s*(from|import)\s+oracle\b)|(liboracle)|(libgravidy_ref)|(hermite_oracle)", re.M)
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h", ".mk")) or f == "Makefile":
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert not bad.search(txt), f"{os.path.join(dp, f)} reaches into oracle/"
