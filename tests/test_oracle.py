"""CPU tests: the oracle (C restatement, oracle/lssvm_oracle.c) against the golden vectors produced by the unmodified
reference (tests/golden/make_golden.py) and — where oracle/_ref is present — against the reference build itself."""
import os
import hashlib

import numpy as np
import pytest

from conftest import nrm_rel
from oracle import port, ref
from plssvm_sparse_fp22_b200 import datagen

KERNELS = ["linear", "polynomial", "rbf"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("kernel", KERNELS)
def test_golden_predict_fixture(golden, kernel):
    """Reference tests/backends/generic_tests.hpp:229-258: the 500x200 models must reproduce the golden labels exactly
    and label * decision > 0."""
    g = golden("predict_500x200.npz")
    degree, gamma, coef0, rho = g[f"{kernel}/params"]
    dec = port.predict(g[f"{kernel}/X"], g[f"{kernel}/alpha"], rho, g["points"], kernel, int(degree), gamma, coef0)
    labels = port.sign(dec)
    assert np.array_equal(labels, g["expected_labels"])
    assert np.all(g["point_labels"] * dec > 0)
    assert nrm_rel(g[f"{kernel}/decision"], dec) < 1e-12


@pytest.mark.parametrize("name,dt", [("f64", np.float64), ("f32", np.float32)])
@pytest.mark.parametrize("kernel", KERNELS)
def test_kernel_q_matvec_known_answers(golden, kernel, name, dt):
    """Reference procedure tests/backends/generic_tests.hpp:120-227 (q and device-kernel checks, d in U(1,2), add = +-1);
    tolerance = the reference's own: |a-b| < 128 eps (|a|+|b|) (tests/utility.hpp:117-136)."""
    g = golden("kernel_kat.npz")
    X, _ = datagen.make_dense(193, 57, seed=11, problem="planes")
    assert sha(X) == str(g["sha_X"]), "datagen drifted from the committed golden inputs"
    X = X.astype(dt)
    d = g["d"].astype(dt)
    degree, gamma, coef0, cost = 3, 1.0 / X.shape[1], 0.5 if kernel == "polynomial" else 0.0, 2.0
    eps = 128 * np.finfo(dt).eps

    def near(a, b, scale=1.0):
        a = np.asarray(a, np.float64)
        b = np.asarray(b, np.float64)
        return np.all(np.abs(a - b) <= np.maximum(scale * eps * (np.abs(a) + np.abs(b)), np.finfo(dt).tiny))

    assert near(port.kernel_function(X[0], X[1], kernel, degree, gamma, coef0), g[f"{name}/{kernel}/k01"])
    q = port.generate_q(X, kernel, degree, gamma, coef0)
    assert near(q, g[f"{name}/{kernel}/q"])
    qa = float(g[f"{name}/{kernel}/QA_cost"])
    for add in (1.0, -1.0):
        want = g[f"{name}/{kernel}/matvec{int(add):+d}"]
        got = port.matvec(X, g[f"{name}/{kernel}/q"], d, qa, add, kernel, degree, gamma, coef0, cost)
        # sums of 192 terms: allow the rounding of the accumulation order (reference is built with -ffast-math)
        assert near(got, want, scale=16.0), (kernel, name, add, np.max(np.abs(got - want)))
        got_rows = port.matvec(X, g[f"{name}/{kernel}/q"], d, qa, add, kernel, degree, gamma, coef0, cost, rows=(0, X.shape[0] - 1))
        assert near(got_rows, want, scale=16.0)


@pytest.mark.parametrize("kernel", KERNELS)
def test_learn_matches_reference_fixture_f64(golden, kernel):
    """CG end to end on the reference's own 500x200 training fixture: identical iteration count, alpha/rho within the
    reference's measured 1-thread-vs-8-thread spread at the BASELINE epsilon 1e-3."""
    g = golden("train_500x200.npz")
    X, y = g["X"], g["y"]
    alpha, rho, it = port.learn(X, y, kernel, eps=1e-3)
    key = f"f64/{kernel}/0.001"
    assert it == int(g[key + "/iters"])
    assert nrm_rel(g[key + "/alpha"], alpha) < max(1e-10, 10 * float(g[key + "/spread_alpha"]))
    rho_ref = float(g[key + "/rho"])
    assert abs(rho - rho_ref) <= max(1e-8, 10 * float(g[key + "/spread_rho"])) * abs(rho_ref) + 1e-12


@pytest.mark.parametrize("case", ["blobs_257x33_s7", "planes_384x64_s8"])
@pytest.mark.parametrize("kernel", KERNELS)
def test_learn_synth_f64(golden, case, kernel):
    g = golden("cg_synth.npz")
    problem, shape, seed = case.split("_")
    n, F = map(int, shape.split("x"))
    X, y = datagen.make_dense(n, F, seed=int(seed[1:]), problem=problem)
    assert sha(X) == str(g[case + "/sha_X"])
    alpha, rho, it = port.learn(X, y, kernel, eps=1e-3, use_rows=True)
    key = f"{case}/f64/{kernel}/0.001"
    assert it == int(g[key + "/iters"])
    assert nrm_rel(g[key + "/alpha"], alpha) < max(1e-10, 10 * float(g[key + "/spread_alpha"]))


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", KERNELS)
def test_port_vs_reference_build(kernel, dt):
    X, y = datagen.make_dense(150, 24, seed=3, problem="planes", dtype=dt)
    q_ref = ref.generate_q(X, kernel)
    q = port.generate_q(X, kernel)
    tol = 64 * np.finfo(dt).eps
    assert np.max(np.abs(q - q_ref) / np.maximum(np.abs(q_ref), 1e-30)) < tol
    d = np.random.default_rng(0).uniform(1, 2, 149).astype(dt)
    qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / 24) + 1.0
    r1 = ref.matvec(X, q_ref, d, qa, 1.0, kernel)
    r2 = port.matvec(X, q_ref, d, qa, 1.0, kernel)
    assert nrm_rel(r1, r2) < 256 * np.finfo(dt).eps
    ref.set_num_threads(1)
    a1, rho1, it1, _ = ref.learn(X, y, kernel, eps=1e-3)
    ref.set_num_threads(8)
    a8, rho8, it8, _ = ref.learn(X, y, kernel, eps=1e-3)
    a2, rho2, it2 = port.learn(X, y, kernel, eps=1e-3)
    assert it1 == it2
    # the reference's own thread-count spread is the noise floor; alpha_n = -sum(alpha) cancels heavily
    floor = nrm_rel(a1, a8)
    assert nrm_rel(a1, a2) < max(1e-9 if dt == np.float64 else 1e-3, 20 * floor), (nrm_rel(a1, a2), floor)


def test_sparse_restatement_matches_dense():
    rp, ci, v, y = datagen.make_sparse_csr(120, 300, 0.05, seed=9)
    X = datagen.csr_to_dense(rp, ci, v, 300)
    rp2, ci2, v2 = datagen.dense_to_csr(X)
    assert np.array_equal(rp, rp2) and np.array_equal(ci, ci2) and np.array_equal(v, v2)
    d = np.random.default_rng(1).uniform(1, 2, 119)
    for kernel in KERNELS:
        q = port.generate_q(X, kernel)
        qs = port.csr_generate_q(rp, ci, v, 300, kernel)
        assert nrm_rel(q, qs) < 1e-13
        a = port.matvec(X, q, d, 2.0, 1.0, kernel, rows=(0, 119))
        b = port.csr_matvec(rp, ci, v, 300, q, d, 2.0, 1.0, kernel)
        assert nrm_rel(a, b) < 1e-12


def test_fp22_codec_known_answers():
    """FP22 is defined by this build (the reference has none): 1 sign | 8 exp | 13 mantissa, RNE."""
    kat = [(0.0, 0x000000), (1.0, 0x3F800000 >> 10), (-2.0, 0xC0000000 >> 10), (0.5, 0x3F000000 >> 10),
           (1.0 + 2.0 ** -13, (0x3F800000 >> 10) + 1),           # exactly representable: 13th mantissa bit
           (1.0 + 2.0 ** -14, 0x3F800000 >> 10),                  # tie -> even (down)
           (1.0 + 3 * 2.0 ** -14, (0x3F800000 >> 10) + 2),        # tie -> even (up)
           (float(np.float32(3.4028235e38)), 0x7F800000 >> 10)]   # max float rounds up to +inf
    for x, code in kat:
        assert port.fp22_encode(x) == code, (x, hex(port.fp22_encode(x)), hex(code))
        assert int(datagen.fp22_encode(np.array([x], np.float32))[0]) == code
    rng = np.random.default_rng(0)
    x = rng.uniform(-1, 1, 10000).astype(np.float32)
    r = datagen.fp22_round(x)
    assert np.all(np.abs(r - x) <= np.abs(x) * 2.0 ** -14 + 1e-45)
    assert np.array_equal(datagen.fp22_round(r), r)  # idempotent
    for xv in x[:200]:
        assert port.fp22_round(float(xv)) == float(datagen.fp22_round(np.array([xv], np.float32))[0])


def test_reference_cuda_backend_library_links():
    """oracle/_ref/liboracle_ref_cuda.so (the reference's CUDA backend for sm_100) must load without unresolved
    symbols and export the two driver entry points — checked on the CPU so that a link error is not first seen on the GPU box."""
    import ctypes
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/liboracle_ref_cuda.so not built (make -C oracle ref_cuda)")
    L = ctypes.CDLL(ref_cuda.lib_path(), mode=os.RTLD_NOW)
    assert L.orefcuda_solver_cg and L.orefcuda_matvec_ms
