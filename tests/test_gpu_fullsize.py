"""Parity at BASELINE.json's full single-GPU sizes through size-independent properties and oracle spot checks.

The oracle (O(m^2 F) on the host) cannot run configs[1] or the sparse configurations in seconds, so at these sizes the implicit
matrix-vector product is checked by
  * rows of the reference formula (src/plssvm/backends/OpenMP/svm_kernel.cpp:31-36) evaluated in double with numpy / scipy for a
    handful of rows i:  (Abar d)_i = sum_j (k(x_i, x_j) + QA_cost - q_i - q_j) d_j + d_i / C,
  * agreement of independent device paths (tcgen05 vs SIMT; cached S vs S recomputed on the fly),
  * symmetry  d1 . (Abar d2) == d2 . (Abar d1)  and linearity  Abar (a d1 + b d2) == a Abar d1 + b Abar d2.
"""
import os

import numpy as np
import pytest

from plssvm_sparse_fp22_b200 import capi, datagen

pytestmark = pytest.mark.gpu


def nrm_rel(want, got):
    want = np.asarray(want, np.float64)
    return float(np.linalg.norm(np.asarray(got, np.float64) - want) / max(np.linalg.norm(want), 1e-300))


def kernel_rows(kernel, G, ni, nj, degree, gamma, coef0):
    if kernel == "rbf":
        return np.exp(-gamma * np.maximum(ni[:, None] + nj[None, :] - 2.0 * G, 0.0))
    if kernel == "polynomial":
        return (gamma * G + coef0) ** degree
    return G


def abar_rows_dense(X, rows, d, q, qa, cost, kernel, degree, gamma, coef0):
    X64 = X.astype(np.float64)
    m = X.shape[0] - 1
    nrm = np.einsum("ij,ij->i", X64, X64)
    G = X64[rows] @ X64[:m].T
    K = kernel_rows(kernel, G, nrm[rows], nrm[:m], degree, gamma, coef0)
    d64, q64 = d.astype(np.float64), q.astype(np.float64)
    return (K + qa - q64[rows][:, None] - q64[None, :]) @ d64 + d64[rows] / cost


def test_dense_configs1_full_size_properties():
    n, F = 65536, 1024
    X, _ = datagen.make_dense(n, F, seed=43, problem="blobs", dtype=np.float32)
    rng = np.random.default_rng(1)
    d1 = rng.standard_normal(n - 1).astype(np.float32)
    d2 = rng.uniform(-1, 2, n - 1).astype(np.float32)
    gamma = 1.0 / F
    with capi.Handle(np.float32, "rbf", 3, gamma, 0.0, 1.0) as h:
        h.set_option(capi.OPT_DENSE_CACHE, capi.CACHE_OFF)  # Gram tiles recomputed on the tensor cores in every matvec
        h.set_data(X)
        q = h.generate_q()
        qa = h.kxx_last() + 1.0
        h.set_q(q, qa)
        y1 = h.matvec(d1)
        assert h.stats().matvec_path == capi.PATH_TC and h.stats().dense_cache_bytes == 0
        y2 = h.matvec(d2)
        y3 = h.matvec((2.0 * d1 - 3.0 * d2).astype(np.float32))
        assert np.array_equal(y1, h.matvec(d1))  # deterministic reductions: bit-reproducible
        h.set_option(capi.OPT_MATVEC_PATH, capi.PATH_SIMT)
        y1_simt = h.matvec(d1)
        assert h.stats().matvec_path == capi.PATH_SIMT
        h.set_option(capi.OPT_MATVEC_PATH, capi.PATH_TC)
        h.set_option(capi.OPT_TC_MODE, capi.TC_F16X3)
        y1_f16 = h.matvec(d1)
        h.set_option(capi.OPT_TC_MODE, capi.TC_FP8X)
        h.set_option(capi.OPT_DENSE_CACHE, capi.CACHE_REQUIRED)  # kernel values stored once (8.6 GB), matvec = sweep over them
        y1_store = h.matvec(d1)  # the STORE pass of the Gram kernel serves this matvec itself
        assert h.stats().dense_cache_bytes > 8e9
        y1_cached = h.matvec(d1)  # sweep over the stored tiles
        assert np.array_equal(y1_cached, h.matvec(d1))
        assert np.array_equal(y1_store, y1)  # the STORE pass keeps the recompute mode's work split and order of sums
    # q against the formula
    X64 = X.astype(np.float64)
    nrm = np.einsum("ij,ij->i", X64, X64)
    q_ref = np.exp(-gamma * np.maximum(nrm[:-1] + nrm[-1] - 2.0 * (X64[:-1] @ X64[-1]), 0.0))
    assert nrm_rel(q_ref, q) < 1e-5
    # independent device paths
    assert nrm_rel(y1_simt, y1) < 2e-5, nrm_rel(y1_simt, y1)
    assert nrm_rel(y1_f16, y1) < 2e-5, nrm_rel(y1_f16, y1)
    assert nrm_rel(y1_cached, y1) < 2e-6, nrm_rel(y1_cached, y1)  # same kernel values, different summation order
    # rows of the reference formula in double
    rows = np.array([0, 1, 127, 128, 255, 256, 4097, 32767, 32768, 50001, 65407, 65408, 65533, 65534])
    want = abar_rows_dense(X, rows, d1, q, qa, 1.0, "rbf", 3, gamma, 0.0)
    scale = np.abs(y1.astype(np.float64)).max()
    assert np.max(np.abs(y1[rows] - want)) < 2e-5 * scale, np.max(np.abs(y1[rows] - want)) / scale
    # symmetry and linearity
    a, b = float(np.dot(d1.astype(np.float64), y2)), float(np.dot(d2.astype(np.float64), y1))
    assert abs(a - b) <= 2e-5 * max(abs(a), abs(b)), (a, b)
    assert nrm_rel(2.0 * y1.astype(np.float64) - 3.0 * y2.astype(np.float64), y3) < 2e-5


def abar_rows_csr(rp, ci, v, F, rows, d, q, qa, cost, kernel, degree, gamma, coef0):
    import scipy.sparse as sps

    n = rp.shape[0] - 1
    m = n - 1
    X = sps.csr_matrix((v.astype(np.float64), ci, rp), shape=(n, F))
    nrm = np.asarray(X.multiply(X).sum(axis=1)).ravel()
    G = (X[rows] @ X[:m].T).toarray()
    K = kernel_rows(kernel, G, nrm[rows], nrm[:m], degree, gamma, coef0)
    d64, q64 = d.astype(np.float64), q.astype(np.float64)
    return (K + qa - q64[rows][:, None] - q64[None, :]) @ d64 + d64[rows] / cost


@pytest.mark.parametrize("case", ["rbf_f64_131072x100k", "poly_fp22_65536x20k"])
def test_sparse_full_size_properties(case):
    if case.startswith("rbf"):
        n, F, density, kernel, dt, coef0 = 131072, 100000, 1e-3, "rbf", np.float64, 0.0
    else:
        n, F, density, kernel, dt, coef0 = 65536, 20000, 1e-2, "polynomial", np.float32, 0.5
    rp, ci, v, _ = datagen.make_sparse_csr(n, F, density, seed=45, dtype=dt)
    fp22 = case.startswith("poly")
    if fp22:
        v = datagen.fp22_round(v)
        prp, packed = datagen.pack_fp22_csr(rp, ci, v)
    rng = np.random.default_rng(2)
    d1 = rng.standard_normal(n - 1).astype(dt)
    d2 = rng.uniform(-1, 2, n - 1).astype(dt)
    gamma = 1.0 / F
    tol = 1e-11 if dt == np.float64 else 2e-5
    out = {}
    for cache in (capi.CACHE_REQUIRED, capi.CACHE_OFF):
        with capi.Handle(dt, kernel, 3, gamma, coef0, 1.0) as h:
            h.set_option(capi.OPT_SPARSE_CACHE, cache)
            if fp22:
                h.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
            else:
                h.set_data_csr(rp, ci, v, F)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            h.set_q(q, qa)
            y1 = h.matvec(d1)
            assert np.array_equal(y1, h.matvec(d1))  # integer accumulation + fixed-order reductions: bit-reproducible
            if cache == capi.CACHE_REQUIRED:
                y2 = h.matvec(d2)
                y3 = h.matvec((2.0 * d1 - 3.0 * d2).astype(dt))
                assert h.stats().sparse_cache_entries > 0
            out[cache] = y1
    y1 = out[capi.CACHE_REQUIRED]
    assert nrm_rel(out[capi.CACHE_OFF], y1) < tol
    rows = np.array([0, 1, 8191, 8192, 40000, n - 3, n - 2])
    want = abar_rows_csr(rp, ci, v, F, rows, d1, q, qa, 1.0, kernel, 3, gamma, coef0)
    scale = np.abs(y1.astype(np.float64)).max()
    assert np.max(np.abs(y1[rows] - want)) < tol * scale, np.max(np.abs(y1[rows] - want)) / scale
    a, b = float(np.dot(d1.astype(np.float64), y2)), float(np.dot(d2.astype(np.float64), y1))
    assert abs(a - b) <= 10 * tol * max(abs(a), abs(b)), (a, b)
    assert nrm_rel(2.0 * y1.astype(np.float64) - 3.0 * y2.astype(np.float64), y3) < 10 * tol


@pytest.mark.parametrize("problem", ["blobs", "planes"])
def test_configs0_linear_double_against_the_reference(problem):
    """BASELINE.json configs[0] at its exact size (5000 points x 256 dense features, double, linear, eps = 1e-3): the reference's
    own OpenMP learn() on the host cores against b200svm_learn — identical CG iteration count, alpha within 1e-10 (rho within its cancellation noise),
    bit-identical labels on the training points."""
    from oracle import port, ref

    n, F = 5000, 256
    X, y = datagen.make_dense(n, F, seed=42, problem=problem, dtype=np.float64)
    if ref.available():  # the unmodified reference, one thread (bit-deterministic); ~1-2 minutes on the host
        ref.set_num_threads(1)
        a_o, rho_o, it_o, _ = ref.learn(X, y, "linear", eps=1e-3)
        ref.set_num_threads(os.cpu_count() or 1)
    else:
        a_o, rho_o, it_o = port.learn(X, y, "linear", eps=1e-3, use_rows=True)
    with capi.Handle(np.float64, "linear", gamma=1.0 / F) as h:
        h.set_data(X)
        alpha, rho, it = h.learn(y, eps=1e-3)
        dec = h.predict(alpha, rho, X)
    assert it == it_o
    assert nrm_rel(a_o, alpha) < 1e-10, nrm_rel(a_o, alpha)
    # rho = y_n + QA_cost sum(alpha) - q.alpha cancels sums of ~5000 terms: the reference differs from itself by
    # 1e-7 relative between thread counts (SURVEY.md §8c noise floor), so rho is held to 1e-6 while alpha is held to 1e-10
    assert abs(rho - rho_o) <= 1e-6 * max(abs(rho_o), 1.0), (rho, rho_o)
    dec_o = port.predict(X, a_o, rho_o, X, "linear")
    safe = np.abs(dec_o) > 1e-9 * np.abs(dec_o).max()
    assert np.array_equal(np.sign(dec[safe]), np.sign(dec_o[safe]))


def test_configs1_full_size_cg_iteration_against_the_reference_golden(golden):
    """BASELINE.json configs[1] at FULL size (65536 x 1024 float rbf): one CG iteration — two applications of the implicit matrix and
    the vector algebra between them — against the UNMODIFIED reference's solver_CG on the same data (tests/golden/cg_cfg1_full.npz,
    written by tests/golden/make_golden_cfg1.py from oracle/_ref: 18 minutes on 8 host cores).  Both the cached-matrix mode (STORE pass
    + sweep) and the recompute mode of the tcgen05 path are held to it."""
    g = golden("cg_cfg1_full.npz")
    n, F = int(g["n"]), int(g["F"])
    X, y = datagen.make_dense(n, F, seed=int(g["seed"]), problem="blobs", dtype=np.float32)
    b = (y[:-1] - y[-1]).astype(np.float32)
    for cache in (capi.CACHE_AUTO, capi.CACHE_OFF):
        with capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0) as h:
            h.set_option(capi.OPT_DENSE_CACHE, cache)
            h.set_data(X)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            assert nrm_rel(g["q_head"], q[:4096]) < 2e-6 and abs(float(q.astype(np.float64).sum()) - float(g["q_sum"])) < 2e-6 * abs(float(g["q_sum"]))
            assert abs(qa - float(g["QA_cost"])) < 1e-6
            alpha, it, delta = h.solver_cg(b, None, qa, 1, 1e-3)
            assert h.stats().matvec_path == capi.PATH_TC
        assert it == int(g["iterations"]) == 1
        err = nrm_rel(g["alpha"], alpha)
        if "alpha_f64" in g.files:
            # the one float policy (oracle/policy.py): against the reference's DOUBLE run of the same iteration, ours must be within
            # FLOAT_K x the error of the reference's own float run (floor 1e-5)
            from oracle import policy
            err_ref = nrm_rel(g["alpha_f64"], g["alpha"])
            err_ours = nrm_rel(g["alpha_f64"], alpha)
            assert err_ours <= policy.float_alpha_tolerance(err_ref, 0.0), (cache, err_ours, err_ref)
        # against the reference's float run itself: its scalars (delta, d.Ad) are sequential float sums over 65535 terms under -ffast-math
        # (operators.hpp:113-130), which alone moves alpha_cd by ~1e-4 relative; measured distance on B200: 1.2e-4
        assert err < 5e-4, (cache, err)
