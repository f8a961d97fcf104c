"""GPU parity tests (run with -m gpu on a B200): every call goes through the C ABI (libb200svm.so via ctypes) and is
compared with the oracle (oracle/lssvm_oracle.c, pinned to the reference by tests/test_oracle.py) on the same seeded
inputs, and with the golden vectors the unmodified reference produced (tests/golden/).

Tolerances (floating point path, stated per BASELINE.json north_star and SURVEY.md F4/H2):
  * q, matvec: the reference's own test tolerance 128 eps (|a|+|b|) per element (tests/utility.hpp:117-136), with a factor for
    the different (tree/tile) summation order; float additionally norm-wise.
  * CG: identical iteration count; alpha norm-wise <= 1e-10 (double) at the BASELINE epsilon 1e-3, or 10x the reference's own
    1-vs-8-thread spread where that is larger; float against 10x the reference's measured float spread (its float noise floor
    is 1e-5..1e-3, tests/golden/make_golden.py output) and never worse than the f64 truth distance of the reference float.
  * rho: cancellation-aware, tol * (|y_n| + |QA_cost sum(alpha)| + |q.alpha|).
  * predicted labels: bit-identical.
"""
import numpy as np
import pytest

from conftest import nrm_rel
from oracle import policy, port
from plssvm_sparse_fp22_b200 import capi, datagen

pytestmark = pytest.mark.gpu
KERNELS = ["linear", "polynomial", "rbf"]
DTYPES = [np.float64, np.float32]


def near(a, b, dt, scale):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    eps = 128 * np.finfo(dt).eps * scale
    return bool(np.all(np.abs(a - b) <= np.maximum(eps * (np.abs(a) + np.abs(b)), np.finfo(dt).tiny)))


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kernel", KERNELS)
def test_generate_q_and_matvec_kat(golden, kernel, dt):
    g = golden("kernel_kat.npz")
    name = "f64" if dt == np.float64 else "f32"
    X, _ = datagen.make_dense(193, 57, seed=11, problem="planes")
    X = X.astype(dt)
    degree, gamma, coef0, cost = 3, 1.0 / X.shape[1], 0.5 if kernel == "polynomial" else 0.0, 2.0
    with capi.Handle(dt, kernel, degree, gamma, coef0, cost) as h:
        h.set_data(X)
        q = h.generate_q()
        assert near(q, g[f"{name}/{kernel}/q"], dt, 8.0), np.max(np.abs(q - g[f"{name}/{kernel}/q"]))
        qa = float(g[f"{name}/{kernel}/QA_cost"])
        kxx = h.kxx_last()
        assert abs((kxx + 1.0 / cost) - qa) <= 1e3 * np.finfo(dt).eps * abs(qa)
        h.set_q(g[f"{name}/{kernel}/q"], qa)
        d = g["d"].astype(dt)
        for add in (1.0, -1.0):
            want = g[f"{name}/{kernel}/matvec{int(add):+d}"]
            got = h.matvec(d, add)
            assert nrm_rel(want, got) < (1e-13 if dt == np.float64 else 2e-6), (kernel, name, add, nrm_rel(want, got))
            # accumulate into a non-zero ret (ret += add * A d)
            base = np.linspace(-1, 1, want.shape[0]).astype(dt)
            got2 = h.matvec(d, add, ret=base)
            assert nrm_rel(want + base, got2) < (1e-13 if dt == np.float64 else 2e-6)


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("shape", [(2, 1), (3, 5), (65, 17), (129, 200), (700, 130)])
def test_matvec_vs_oracle_shapes(kernel, dt, shape):
    """Ragged / tiny / tile-edge sizes against the oracle's full-row operator."""
    n, F = shape
    X, _ = datagen.make_dense(n, F, seed=100 + n, problem="blobs", dtype=dt)
    rng = np.random.default_rng(n)
    d = rng.uniform(-1, 2, n - 1).astype(dt)
    with capi.Handle(dt, kernel, 3, 1.0 / F, 0.25 if kernel == "polynomial" else 0.0, 1.5) as h:
        h.set_data(X)
        q = h.generate_q()
        q_o = port.generate_q(X, kernel, 3, 1.0 / F, 0.25 if kernel == "polynomial" else 0.0)
        assert nrm_rel(q_o, q) < (1e-13 if dt == np.float64 else 5e-6)
        qa = h.kxx_last() + 1.0 / 1.5
        h.set_q(q_o, qa)
        got = h.matvec(d, 1.0)
        want = port.matvec(X.astype(np.float64), q_o.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 3, 1.0 / F,
                           0.25 if kernel == "polynomial" else 0.0, 1.5, rows=(0, n - 1))
        assert nrm_rel(want, got) < (1e-12 if dt == np.float64 else 5e-5), nrm_rel(want, got)


@pytest.mark.parametrize("kernel", KERNELS)
def test_learn_reference_fixture_f64(golden, kernel):
    """The reference's own 500x200 training fixture, double, epsilon 1e-3 (BASELINE config-1 setting):
    identical CG iteration count, alpha <= 1e-10 norm-wise (or 10x the reference's own thread spread), rho cancellation-aware."""
    g = golden("train_500x200.npz")
    X, y = g["X"], g["y"]
    key = f"f64/{kernel}/0.001"
    with capi.Handle(np.float64, kernel) as h:
        h.set_data(X)
        alpha, rho, it = h.learn(y, eps=1e-3)
    assert it == int(g[key + "/iters"])
    tol = max(1e-10, 10 * float(g[key + "/spread_alpha"]))
    assert nrm_rel(g[key + "/alpha"], alpha) < tol, (nrm_rel(g[key + "/alpha"], alpha), tol)
    a_ref = g[key + "/alpha"]
    q = port.generate_q(X, kernel)
    qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / X.shape[1]) + 1.0
    scale = abs(y[-1]) + abs(qa * a_ref[:-1].sum()) + abs(q @ a_ref[:-1])
    assert abs(rho - float(g[key + "/rho"])) <= 1e-10 * scale + 10 * float(g[key + "/spread_rho"]) * abs(float(g[key + "/rho"]))


@pytest.mark.parametrize("eps", [1e-3, 1e-6])
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("case", ["blobs_257x33_s7", "planes_384x64_s8", "blobs_1000x256_s43"])
def test_learn_synth_golden(golden, case, kernel, eps):
    g = golden("cg_synth.npz")
    problem, shape, seed = case.split("_")
    n, F = map(int, shape.split("x"))
    X, y = datagen.make_dense(n, F, seed=int(seed[1:]), problem=problem)
    for name, dt in (("f64", np.float64), ("f32", np.float32)):
        key = f"{case}/{name}/{kernel}/{eps:g}"
        with capi.Handle(dt, kernel) as h:
            h.set_data(X.astype(dt))
            alpha, rho, it = h.learn(y.astype(dt), eps=eps)
        it_ref, it_ref8 = int(g[key + "/iters"]), int(g[key + "/iters8"])
        spread = float(g[key + "/spread_alpha"])
        if dt == np.float64 and eps == 1e-3:
            assert it == it_ref
            assert nrm_rel(g[key + "/alpha"], alpha) < max(1e-10, 10 * spread)
        elif dt == np.float64:
            # tight epsilon: the reference itself is not reproducible across thread counts
            assert abs(it - it_ref) <= max(1, abs(it_ref - it_ref8))
            # CG stops at |r| <= eps |r0|: two rounding-different runs may legitimately differ by O(cond * eps)
            assert nrm_rel(g[key + "/alpha"], alpha) < max(1e-8, 20 * spread, 100 * eps)
        else:
            # float: compare with the f64 truth; must be at least as close as the reference's float result (x2 slack)
            truth = g[f"{case}/f64/{kernel}/{eps:g}/alpha"]
            err_ref = nrm_rel(truth, g[key + "/alpha"])
            err_ours = nrm_rel(truth, alpha)
            assert abs(it - it_ref) <= max(2, abs(it_ref - it_ref8), it_ref // 4), (it, it_ref)
            assert err_ours <= policy.float_alpha_tolerance(err_ref, spread), (err_ours, err_ref, spread)  # the one float policy (oracle/policy.py)


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kernel", KERNELS)
def test_solver_cg_direct_and_refresh(kernel, dt):
    """solver_CG called directly with imax large and eps tiny so the every-50th-iteration residual refresh
    (src/plssvm/backends/OpenMP/csvm.cpp:130-135) is exercised; compared with the oracle's CG."""
    n, F = 300, 20
    X, y = datagen.make_dense(n, F, seed=77, problem="planes", dtype=dt)
    q = port.generate_q(X, kernel)
    qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / F) + 1.0
    b = (y[:-1] - y[-1]).astype(dt)
    # float: the recursively updated residual underflows to 0 after ~18 iterations (oracle and GPU alike) -> stay below that
    imax = 60 if dt == np.float64 else 8
    a_o, it_o, delta_o = port.solver_cg(X, b, q, qa, imax, 0.0, kernel, use_rows=True)
    with capi.Handle(dt, kernel) as h:
        h.set_data(X)
        a, it, delta = h.solver_cg(b, q, qa, imax, 0.0)
    assert it == it_o == imax
    if dt == np.float64:
        assert nrm_rel(a_o, a) < 1e-6, nrm_rel(a_o, a)  # 60 CG steps amplify rounding-order differences
    else:
        # two float CG runs pushed past convergence (eps = 0) drift apart by rounding order; judge both against the f64 solution
        # of the same 8 iterations: ours must be as close to it as the oracle's float run is (x2 slack, floor 1e-3)
        X64 = X.astype(np.float64)
        a_t, _, _ = port.solver_cg(X64, b.astype(np.float64), q.astype(np.float64), float(qa), imax, 0.0, kernel, use_rows=True)
        assert nrm_rel(a_t, a) < max(2 * nrm_rel(a_t, a_o), 1e-3), (nrm_rel(a_t, a), nrm_rel(a_t, a_o))


@pytest.mark.parametrize("dt", DTYPES)
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("eps,imax", [(1e-3, 64), (1e-8, 120), (0.0, 7)])
def test_cg_lookahead_depth_does_not_change_the_result(kernel, dt, eps, imax):
    """The host queues up to k iterations beyond the one whose delta it has read (B200SVM_OPT_SYNC_EVERY); the device-side stop
    flag must freeze alpha / delta at the converged iteration: bit-identical alpha, same iteration count and delta for every k,
    including the every-50th-iteration residual refresh (csvm.cpp:130-135) and convergence on the last allowed iteration."""
    n, F = 400, 24
    X, y = datagen.make_dense(n, F, seed=5, problem="planes", dtype=dt)
    b = (y[:-1] - y[-1]).astype(dt)
    results = []
    for ahead in (0, 1, 3, 8):
        with capi.Handle(dt, kernel, gamma=1.0 / F) as h:
            h.set_option(capi.OPT_SYNC_EVERY, ahead)
            h.set_data(X)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            results.append(h.solver_cg(b, q, qa, imax, eps))
            # the handle is reusable afterwards (flag cleared): further solves repeat each other bit for bit.  (The very first solve on
            # a data set may differ from them in the last bits when its first matvec assembled the dense kernel-matrix cache.)
            again = h.solver_cg(b, q, qa, imax, eps)
            third = h.solver_cg(b, q, qa, imax, eps)
            assert again[1] == third[1] and np.array_equal(again[0], third[0])
            if h.stats().dense_cache_bytes == 0:
                assert again[1] == results[-1][1] and np.array_equal(again[0], results[-1][0])
            elif eps >= 1e-3 and again[1] == results[-1][1]:  # (runs pushed past convergence drift apart by rounding order)
                assert nrm_rel(results[-1][0], again[0]) < 1e-2
            st = h.stats()
            assert st.matvecs_timed == st.matvecs
    a0, it0, d0 = results[0]
    for a, it, d in results[1:]:
        assert it == it0 and d == d0
        assert np.array_equal(a, a0)
    if dt == np.float64 and eps > 0:
        a_o, it_o, _ = port.solver_cg(X, b, port.generate_q(X, kernel), port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / F) + 1.0, imax, eps, kernel, use_rows=True)
        if eps == 1e-3:
            assert it0 == it_o
        else:
            assert abs(it0 - it_o) <= 2  # at eps = 1e-8 the recursion's rounding decides the last step or two


@pytest.mark.parametrize("kernel", KERNELS)
def test_predict_golden_labels(golden, kernel):
    """Reference golden fixture (tests/backends/generic_tests.hpp:229-258): bit-identical labels, label * decision > 0."""
    g = golden("predict_500x200.npz")
    degree, gamma, coef0, rho = g[f"{kernel}/params"]
    for dt in DTYPES:
        with capi.Handle(dt, kernel, int(degree), gamma if gamma > 0 else 1.0 / 200, coef0) as h:
            h.set_data(g[f"{kernel}/X"].astype(dt))
            dec = h.predict(g[f"{kernel}/alpha"].astype(dt), rho, g["points"].astype(dt))
            labels = h.predict_label(dec)
            assert np.array_equal(labels, g["expected_labels"].astype(dt))
            assert np.all(g["point_labels"] * dec > 0)
            assert nrm_rel(g[f"{kernel}/decision"], dec) < (1e-12 if dt == np.float64 else 1e-4)
            if kernel == "linear":
                w = h.update_w(g[f"{kernel}/alpha"].astype(dt))
                w_o = port.update_w(g[f"{kernel}/X"], g[f"{kernel}/alpha"])
                assert nrm_rel(w_o, w) < (1e-13 if dt == np.float64 else 1e-5)


def test_predict_edge_cases():
    X, y = datagen.make_dense(50, 8, seed=5)
    with capi.Handle(np.float64, "rbf") as h:
        h.set_data(X)
        alpha = np.linspace(-1, 1, 50)
        out = h.predict(alpha, 0.3, np.empty((0, 8)))
        assert out.shape == (0,)
        one = h.predict(alpha, 0.3, X[:1])
        want = port.predict(X, alpha, 0.3, X[:1], "rbf")
        assert abs(one[0] - want[0]) < 1e-12
        # 0 decision -> label -1 (operators::sign)
        assert h.predict_label(np.array([0.0, 1e-300, -0.0]))[0] == -1.0


def test_null_handle_is_rejected():
    L = capi.load()
    assert L.b200svm_set_option(None, capi.OPT_MATVEC_PATH, 1) == 1
    assert b"null" in L.b200svm_last_error(None)


def test_errors_are_loud():
    with capi.Handle(np.float64, "linear", gamma=1.0) as h:
        with pytest.raises(capi.B200Error, match="no data set"):
            h.n, h.F = 3, 2
            h.matvec(np.ones(2))
        with pytest.raises(capi.B200Error, match="Data set is empty"):
            h.L.b200svm_set_data_dense.restype = int
            h._chk(h.L.b200svm_set_data_dense(h.h, np.zeros(1).ctypes.data, 0, 4, 4))


@pytest.mark.parametrize("cache", [capi.CACHE_OFF, capi.CACHE_REQUIRED])
@pytest.mark.parametrize("mode", [capi.TC_F16X3, capi.TC_FP8X])
@pytest.mark.parametrize("kernel", ["rbf", "polynomial"])
@pytest.mark.parametrize("shape", [(2, 3), (130, 64), (257, 33), (513, 100), (1000, 200), (2049, 1024)])
def test_tcgen05_path_vs_simt_and_oracle(kernel, shape, mode, cache):
    """The tcgen05/TMA path (fp16x2-split Gram tiles, symmetric) against the SIMT fp32 path and the f64 oracle on ragged sizes
    (rows not a multiple of the 128x256 tile, features not a multiple of the 64-wide K block).  cache OFF: the Gram tiles are
    recomputed in every matvec; REQUIRED: the kernel values are stored by one pass of the same kernel and swept by k_cached_gemv."""
    n, F = shape
    X, _ = datagen.make_dense(n, F, seed=n + F, problem="blobs", dtype=np.float32)
    d = np.random.default_rng(1).uniform(-1, 2, n - 1).astype(np.float32)
    coef0 = 0.5 if kernel == "polynomial" else 0.0
    outs = {}
    for path in (capi.PATH_SIMT, capi.PATH_TC):
        with capi.Handle(np.float32, kernel, 3, 1.0 / F, coef0, 1.0) as h:
            h.set_option(capi.OPT_MATVEC_PATH, path)
            h.set_option(capi.OPT_TC_MODE, mode)
            h.set_option(capi.OPT_DENSE_CACHE, cache)
            h.set_data(X)
            q = h.generate_q()
            h.set_q(q, h.kxx_last() + 1.0)
            outs[path] = h.matvec(d, 1.0)
            assert h.stats().matvec_path == path
            assert (h.stats().dense_cache_bytes > 0) == (path == capi.PATH_TC and cache == capi.CACHE_REQUIRED)
            outs[(path, "neg")] = h.matvec(d, -1.0, ret=np.ones(n - 1, np.float32))
    X64 = X.astype(np.float64)
    q64 = port.generate_q(X64, kernel, 3, 1.0 / F, coef0)
    qa64 = port.kernel_function(X64[-1], X64[-1], kernel, 3, 1.0 / F, coef0) + 1.0
    want = port.matvec(X64, q64, d.astype(np.float64), qa64, 1.0, kernel, 3, 1.0 / F, coef0, 1.0, rows=(0, n - 1))
    assert nrm_rel(want, outs[capi.PATH_SIMT]) < 2e-6
    assert nrm_rel(want, outs[capi.PATH_TC]) < 5e-6, nrm_rel(want, outs[capi.PATH_TC])
    assert nrm_rel(1.0 - want, outs[(capi.PATH_TC, "neg")]) < 5e-6


def test_tcgen05_path_is_deterministic():
    """Bit-reproducible for a given call sequence.  With the kernel-matrix cache the first matvec after set_data is served by the
    assembling (STORE) pass of the Gram kernel — bit for bit the recompute mode's answer — and the later ones by sweeps over the
    stored tiles: same kernel values, another fixed order of the sums."""
    n, F = 1500, 96
    X, _ = datagen.make_dense(n, F, seed=3, dtype=np.float32)
    d = np.random.default_rng(2).standard_normal(n - 1).astype(np.float32)
    for kernel in ("rbf", "polynomial"):
        first, later = {}, {}
        for cache in (capi.CACHE_OFF, capi.CACHE_REQUIRED):
            with capi.Handle(np.float32, kernel, gamma=1.0 / F, coef0=0.25) as h:
                h.set_option(capi.OPT_MATVEC_PATH, capi.PATH_TC)
                h.set_option(capi.OPT_DENSE_CACHE, cache)
                h.set_data(X)
                q = h.generate_q()
                h.set_q(q, h.kxx_last() + 1.0)
                first[cache] = h.matvec(d)
                later[cache] = h.matvec(d)
                for _ in range(3):
                    assert np.array_equal(later[cache], h.matvec(d))
                assert (h.stats().dense_cache_bytes > 0) == (cache == capi.CACHE_REQUIRED)
        assert np.array_equal(first[capi.CACHE_OFF], later[capi.CACHE_OFF])
        assert np.array_equal(first[capi.CACHE_REQUIRED], first[capi.CACHE_OFF])
        assert nrm_rel(later[capi.CACHE_OFF], later[capi.CACHE_REQUIRED]) < 1e-5


def test_multi_gpu_row_and_tile_partition():
    """N >= 2 GPUs only (skipped on a 1-GPU box): tests/multi_gpu_check.py under torchrun — NCCL allreduce of the per-rank partial
    K d, identical iteration counts and bit-identical alpha on every rank."""
    import subprocess
    import sys
    import torch
    ngpu = torch.cuda.device_count()
    if ngpu < 2:
        pytest.skip("needs >= 2 GPUs")
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={min(ngpu, 8)}", "--master-addr", "127.0.0.1",
                        "--master-port", "29541", os.path.join(root, "tests", "multi_gpu_check.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("kernel", ["linear", "polynomial", "rbf"])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_reference_cuda_backend_agrees(kernel, dt):
    """the reference's own CUDA kernels (unmodified, built for sm_100: oracle/_ref/liboracle_ref_cuda.so) and
    libb200svm compute the same ret += add * Abar d on the same device (src/plssvm/backends/CUDA/svm_kernel.cu vs csrc/)."""
    from oracle import ref_cuda
    if not ref_cuda.available():
        pytest.skip("oracle/_ref/liboracle_ref_cuda.so not built")
    n, F = 700, 90
    X, _ = datagen.make_dense(n, F, seed=5, problem="planes", dtype=dt)
    rng = np.random.default_rng(2)
    d = rng.uniform(-1, 2, n - 1).astype(dt)
    with capi.Handle(dt, kernel, 3, 1.0 / F, 0.0, 1.5) as h:
        h.set_data(X)
        q = h.generate_q()
        qa = h.kxx_last() + 1.0 / 1.5
        h.set_q(q, qa)
        ours = h.matvec(d, 1.0)
    theirs, _, _ = ref_cuda.matvec_ms(X, q, d, qa, 1.0, kernel, cost=1.5, reps=1)
    rel = float(np.linalg.norm(ours.astype(np.float64) - theirs) / np.linalg.norm(theirs))
    assert rel < (1e-12 if dt == np.float64 else 3e-5), rel  # (float: the reference kernel is built with --use_fast_math and sums with atomics)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("dt", DTYPES)
def test_torch_device_tensor_api(kernel, dt):
    """fit / predict on torch CUDA tensors (device pointers through the C ABI) == the host-buffer path."""
    import torch
    from plssvm_sparse_fp22_b200.torch_api import LSSVM
    n, F = 600, 40
    X, y = datagen.make_dense(n, F, seed=11, problem="planes", dtype=dt)
    with capi.Handle(dt, kernel) as h:
        h.set_data(X)
        alpha, rho, it = h.learn(y, eps=1e-3)
        dec = h.predict(alpha, rho, X[:100])
    Xt, yt = torch.from_numpy(X).cuda(), torch.from_numpy(y).cuda()
    svm = LSSVM(kernel=kernel).fit(Xt, yt)
    assert svm.iterations == it
    a_t = svm.alpha.cpu().numpy()
    assert nrm_rel(alpha, a_t) < (1e-12 if dt == np.float64 else 1e-4)
    v = svm.decision_function(Xt[:100]).cpu().numpy()
    assert np.allclose(v, dec, rtol=0, atol=(1e-9 if dt == np.float64 else 2e-3) * max(1.0, np.abs(dec).max()))
    lab = svm.predict(Xt[:100]).cpu().numpy()
    assert np.array_equal(lab, capi.Handle.predict_label(v))
    with pytest.raises(TypeError):
        LSSVM(kernel=kernel).fit(torch.from_numpy(X), torch.from_numpy(y))  # CPU tensors: no CPU path
    svm.close()


@pytest.mark.parametrize("kernel", ["polynomial", "rbf"])
@pytest.mark.parametrize("shape", [(130, 20), (700, 33), (1500, 96), (2600, 40)])
def test_dense_double_tile_cache(kernel, shape):
    """Dense double polynomial / rbf: the kernel matrix cached as lower-triangle 128 x 256 tiles (STORE pass of the SIMT Gram kernel +
    symmetric sweep) against the implicit SIMT path on the same handle data and against the oracle."""
    n, F = shape
    X, y = datagen.make_dense(n, F, seed=n, problem="planes", dtype=np.float64)
    d = np.random.default_rng(n).uniform(-1, 2, n - 1)
    res = {}
    for cache in (capi.CACHE_REQUIRED, capi.CACHE_OFF):
        with capi.Handle(np.float64, kernel, gamma=1.0 / F) as h:
            h.set_option(capi.OPT_DENSE_CACHE, cache)
            h.set_data(X)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            h.set_q(q, qa)
            mv = h.matvec(d, 1.0)
            assert np.array_equal(mv, h.matvec(d, 1.0))
            st = h.stats()
            assert (st.dense_cache_bytes > 0) == (cache == capi.CACHE_REQUIRED)
            alpha, rho, it = h.learn(y, eps=1e-3)
            res[cache] = (mv, alpha, it, q, qa)
    mv, alpha, it, q, qa = res[capi.CACHE_REQUIRED]
    want = port.matvec(X, q, d, qa, 1.0, kernel, rows=(0, n - 1))
    assert nrm_rel(want, mv) < 1e-12
    assert nrm_rel(res[capi.CACHE_OFF][0], mv) < 1e-12
    a_o, rho_o, it_o = port.learn(X, y, kernel, eps=1e-3, use_rows=True)
    assert it == it_o == res[capi.CACHE_OFF][2]
    assert nrm_rel(a_o, alpha) < 1e-10
