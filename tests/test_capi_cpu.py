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


def _cuobjdump(*args):
    import shutil
    import subprocess
    exe = shutil.which("cuobjdump") or "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(exe):
        pytest.skip("cuobjdump not available")
    from plssvm_sparse_fp22_b200 import build
    return subprocess.run([exe, *args, build.LIB], capture_output=True, text=True, check=True).stdout


def test_hot_kernels_have_no_spills_and_keep_their_register_budgets():
    """Build-quality gate that needs no GPU: resource usage of the sm_100a code in libb200svm.so.  The hot kernels must not use a
    local-memory stack (spills), and must stay inside the register budget their occupancy was designed for."""
    import re
    out = _cuobjdump("-res-usage")
    usage = {}
    for name, reg, stack in re.findall(r"Function (\S+):\s*\n\s*REG:(\d+) STACK:(\d+)", out):
        usage[name] = (int(reg), int(stack))
    assert usage, out[:500]
    budgets = {  # substring of the mangled name -> registers per thread allowed
        "9k_gram_tc": 204,        # 320 threads, one CTA per SM: 65536 / 320
        "12k_tile_sweep": 128,    # 256 threads, two CTAs per SM
        "14k_sp_cached_mv": 128,
        "10k_kmv_tile": 255,
        "8k_sp_kmv": 255,
    }
    seen = {k: 0 for k in budgets}
    for name, (reg, stack) in usage.items():
        for key, cap in budgets.items():
            if key in name:
                seen[key] += 1
                assert reg <= cap, (name, reg, cap)
                if "ELb0ELb1EE" in name and key == "10k_kmv_tile":
                    assert stack <= 16, (name, stack)  # STORE pass of the double Gram kernel (once per data set): two registers of the exp() epilogue
                elif key != "8k_sp_kmv":
                    assert stack == 0, (name, stack)
    assert all(v > 0 for v in seen.values()), seen


def test_tensor_core_kernel_is_tcgen05_and_tma():
    """The dense Gram kernel must be the tcgen05 / TMA / TMEM program DESIGN.md §3.1 describes (SASS mnemonics from B200_PROFILING.md),
    and the sweep over the cached tiles must use 128-bit streaming loads."""
    sass = _cuobjdump("-sass", "-fun", "_ZN4b2002tc9k_gram_tcILi2ELb1ELb1EEEv14CUtensorMap_stS2_S2_NS0_6paramsE")
    for mnemonic in ("UTCHMMA", "UTCQMMA", "UTMALDG", "LDTM", "SYNCS", "STG.E.EF.128"):
        assert mnemonic in sass, mnemonic
    sweep = _cuobjdump("-sass", "-fun", "_ZN4b2002tc12k_tile_sweepILb1EEEvNS0_12sweep_paramsE")
    assert sweep.count("LDG.E.EF.128") >= 32 and "SHFL.BFLY" in sweep
