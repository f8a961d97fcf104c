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


def test_stats_struct_and_options_match_the_header():
    """capi.Stats mirrors `b200svm_stats` field by field and the OPT_* constants equal the header's #defines."""
    import re

    from plssvm_sparse_fp22_b200 import capi

    text = open(os.path.join(ROOT, "include", "b200svm.h")).read()
    body = re.search(r"typedef struct b200svm_stats \{(.*?)\} b200svm_stats;", text, re.S).group(1)
    fields = re.findall(r"^\s*(int64_t|double)\s+(\w+);", body, re.M)
    ctype = {"int64_t": "c_long", "double": "c_double"}
    got = [(f[1].__name__.replace("c_int64", "c_long").replace("c_longlong", "c_long"), f[0]) for f in capi.Stats._fields_]
    assert got == [(ctype[t], n) for t, n in fields]
    defines = dict((n, int(v)) for n, v in re.findall(r"#define\s+B200SVM_(OPT_\w+)\s+(\d+)", text))
    assert len(defines) >= 11 and len(set(defines.values())) == len(defines)
    for name, value in defines.items():
        assert getattr(capi, name) == value, name


def test_bench_reference_arm_contract(tmp_path):
    """`bench.py --impl reference` prints one JSON line with the contract's keys on rank 0 and nothing on the other ranks."""
    import json
    import subprocess
    import sys

    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--points", "384", "--features", "32", "--steps", "1", "--warmup", "0"]
    env = dict(os.environ, OMP_NUM_THREADS="2")
    r = subprocess.run(cmd, capture_output=True, text=True, env=env, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "cg_iterations_per_sec" and d["unit"] == "iterations/s" and d["value"] > 0
    assert d["higher_is_better"] is True and d["vs_baseline"] is None and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1 and d["cpu_baseline"]["value"] == d["value"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    r1 = subprocess.run(cmd, capture_output=True, text=True, env=dict(env, RANK="1", WORLD_SIZE="2"), timeout=60)
    assert r1.returncode == 0 and r1.stdout.strip() == ""
