"""CPU tests of the boundary: libb200svm.so loads, exports every symbol include/b200svm.h declares, and fails loudly
(no CPU fallback) when no B200 is present.  No compute calls here."""
import ctypes as C
import os
import re

import pytest

from plssvm_sparse_fp22_b200 import build, capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    build.build()
    return capi.load()


def test_header_symbols_all_exported(lib):
    hdr = open(os.path.join(ROOT, "include", "b200svm.h")).read()
    declared = sorted(set(re.findall(r"\b(b200svm_[a-z0-9_]+)\s*\(", hdr)))
    assert declared, "no declarations found"
    assert sorted(capi.SYMBOLS) == declared, "capi.SYMBOLS out of sync with include/b200svm.h"
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in b200svm.h but not exported"


def test_version(lib):
    assert lib.b200svm_version() >= 1000


def test_fails_loudly_without_gpu(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    h = C.c_void_p()
    rc = lib.b200svm_create(C.byref(h), capi.F64, 0, 3, 1.0, 0.0, 1.0, 0)
    assert rc == 2  # B200SVM_ERR_CUDA
    msg = lib.b200svm_last_error(None).decode()
    assert "no CPU fallback" in msg


def test_product_never_imports_oracle():
    """The product path must not route through the oracle (only tests/, smoke() and bench.py's cpu_baseline may)."""
    pkg = os.path.join(ROOT, "plssvm_sparse_fp22_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                txt = open(os.path.join(dirpath, f), errors="ignore").read()
                assert "import oracle" not in txt and "from oracle" not in txt and "lssvm_oracle" not in txt, f
