"""GPU parity tests of the sparse (CSR / FP22-packed) path through the C ABI, against the oracle's CSR restatement
(oracle/lssvm_oracle.c lssvm_csr_*, itself checked against the dense restatement and the reference build in
tests/test_oracle.py) on the same seeded inputs.  The reference stores data dense, so the semantics are the
dense ones applied to rows whose zeros are not stored.

Tolerances: double 1e-10 norm-wise on alpha with identical iteration count (north star); q / matvec 1e-12 (double), 2e-5 (float):
the sparse path evaluates K_ij = u_i u_j (1 + expm1(2 gamma G_ij)) instead of exp(-gamma |x_i - x_j|^2) — same value, different
rounding.  FP22: the oracle is fed decode(encode(x)), so the register-decoded values must match to float tolerance.
"""
import numpy as np
import pytest

from conftest import nrm_rel
from oracle import port
from plssvm_sparse_fp22_b200 import capi, datagen

pytestmark = pytest.mark.gpu
KERNELS = ["linear", "polynomial", "rbf"]


def make(n, F, density, seed, dt, min_nnz=1):
    rp, ci, v, y = datagen.make_sparse_csr(n, F, density, seed=seed, dtype=dt, min_nnz=min_nnz)
    return rp, ci, v, y


@pytest.mark.parametrize("cache", [capi.CACHE_OFF, capi.CACHE_REQUIRED])
@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("shape", [(2, 7, 0.5, 0), (67, 40, 0.2, 128), (300, 1000, 0.01, 128), (1100, 300, 0.05, 256), (700, 5000, 0.004, 0)])
def test_csr_q_and_matvec_vs_oracle(kernel, dt, shape, cache):
    """cache OFF: S recomputed on the fly in every matvec; REQUIRED: S assembled once, matvec = SpMV over the stored entries."""
    n, F, density, W = shape
    rp, ci, v, _ = make(n, F, density, seed=n + F, dt=dt, min_nnz=0 if n > 2 else 1)
    coef0 = 0.5 if kernel == "polynomial" else 0.0
    d = np.random.default_rng(n).uniform(-1, 2, n - 1).astype(dt)
    with capi.Handle(dt, kernel, 3, 1.0 / F, coef0, 1.5) as h:
        if W:
            h.set_option(capi.OPT_SPARSE_BLOCK, W)
        h.set_option(capi.OPT_SPARSE_CACHE, cache)
        h.set_data_csr(rp, ci, v, F)
        q = h.generate_q()
        q_o = port.csr_generate_q(rp, ci, v, F, kernel, 3, 1.0 / F, coef0)
        assert nrm_rel(q_o, q) < (1e-13 if dt == np.float64 else 5e-6)
        X = datagen.csr_to_dense(rp, ci, v, F)
        kxx = port.kernel_function(X[-1], X[-1], kernel, 3, 1.0 / F, coef0)
        assert abs(h.kxx_last() - kxx) <= 1e3 * np.finfo(dt).eps * max(abs(kxx), 1.0)
        qa = float(kxx) + 1.0 / 1.5
        h.set_q(q_o, qa)
        got = h.matvec(d, 1.0)
        assert h.stats().matvec_path == 4
        rp64, ci64, v64 = rp, ci, v.astype(np.float64)
        want = port.csr_matvec(rp64, ci64, v64, F, q_o.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 3, 1.0 / F, coef0, 1.5)
        assert nrm_rel(want, got) < (1e-12 if dt == np.float64 else 2e-5), nrm_rel(want, got)
        base = np.linspace(-1, 1, n - 1).astype(dt)
        got2 = h.matvec(d, -1.0, ret=base)
        assert nrm_rel(base - want, got2) < (1e-12 if dt == np.float64 else 2e-5)
        if kernel != "linear":
            assert (h.stats().sparse_cache_entries > 0) == (cache == capi.CACHE_REQUIRED)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("shape", [(400, 600, 0.02), (1500, 4000, 0.005)])
def test_csr_learn_f64_matches_dense_oracle(kernel, shape):
    """Full learn() on CSR rows == the oracle's learn() on the densified rows: identical CG iteration count, alpha <= 1e-10."""
    n, F, density = shape
    rp, ci, v, y = make(n, F, density, seed=3 * n, dt=np.float64)
    X = datagen.csr_to_dense(rp, ci, v, F)
    a_o, rho_o, it_o = port.learn(X, y, kernel, eps=1e-3, use_rows=True)
    with capi.Handle(np.float64, kernel) as h:
        h.set_data_csr(rp, ci, v, F)
        alpha, rho, it = h.learn(y, eps=1e-3)
    assert it == it_o
    assert nrm_rel(a_o, alpha) < 1e-10, nrm_rel(a_o, alpha)
    q = port.generate_q(X, kernel)
    qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / F) + 1.0
    scale = abs(y[-1]) + abs(qa * a_o[:-1].sum()) + abs(q @ a_o[:-1])
    assert abs(rho - rho_o) <= 1e-10 * scale


@pytest.mark.parametrize("kernel", KERNELS)
def test_csr_matches_dense_device_path(kernel):
    """Same data through set_data_dense and set_data_csr: the two device paths agree (double)."""
    n, F = 500, 300
    rp, ci, v, y = make(n, F, 0.05, seed=9, dt=np.float64)
    X = datagen.csr_to_dense(rp, ci, v, F)
    with capi.Handle(np.float64, kernel) as hd, capi.Handle(np.float64, kernel) as hs:
        hd.set_data(X)
        hs.set_data_csr(rp, ci, v, F)
        ad, rd, itd = hd.learn(y, eps=1e-3)
        as_, rs, its = hs.learn(y, eps=1e-3)
        assert itd == its
        assert nrm_rel(ad, as_) < 1e-9
        # tight epsilon: more iterations amplify the rounding-order difference of the two paths (O(cond * eps))
        ad, rd, itd = hd.learn(y, eps=1e-6)
        as_, rs, its = hs.learn(y, eps=1e-6)
        assert abs(itd - its) <= 1
        assert nrm_rel(ad, as_) < 1e-4


@pytest.mark.parametrize("kernel", KERNELS)
def test_fp22_packed_path(kernel):
    """FP22-packed CSR (decoded in registers) against the oracle fed with decode(encode(x)) — float."""
    n, F = 900, 6000   # gaps > 1023 occur -> padding entries are exercised
    rp, ci, v, y = make(n, F, 0.004, seed=21, dt=np.float32, min_nnz=0)
    prp, packed = datagen.pack_fp22_csr(rp, ci, v)
    assert packed.shape[0] > rp[-1]
    vq = datagen.fp22_round(v)
    coef0 = 0.25 if kernel == "polynomial" else 0.0
    d = np.random.default_rng(5).uniform(-1, 2, n - 1).astype(np.float32)
    with capi.Handle(np.float32, kernel, 3, 1.0 / F, coef0, 1.0) as h:
        h.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
        q = h.generate_q()
        q_o = port.csr_generate_q(rp, ci, vq, F, kernel, 3, 1.0 / F, coef0)
        assert nrm_rel(q_o, q) < 5e-6
        X = datagen.csr_to_dense(rp, ci, vq, F)
        qa = float(port.kernel_function(X[-1], X[-1], kernel, 3, 1.0 / F, coef0)) + 1.0
        h.set_q(q_o, qa)
        got = h.matvec(d, 1.0)
        want = port.csr_matvec(rp, ci, vq.astype(np.float64), F, q_o.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 3, 1.0 / F, coef0, 1.0)
        assert nrm_rel(want, got) < 2e-5, nrm_rel(want, got)
        # the same rows as plain float CSR of the rounded values give the same operator
        with capi.Handle(np.float32, kernel, 3, 1.0 / F, coef0, 1.0) as h2:
            h2.set_option(capi.OPT_SPARSE_BLOCK, 1024)
            h2.set_data_csr(rp, ci, vq, F)
            h2.set_q(q_o, qa)
            assert nrm_rel(h2.matvec(d, 1.0), got) < 2e-6
        # predict with packed points
        alpha = np.random.default_rng(6).standard_normal(n).astype(np.float32) * 0.1
        dec = h.predict_csr(alpha, 0.1, prp, None, packed, value_format=capi.VAL_FP22)
        dec_o = port.csr_predict(rp, ci, vq.astype(np.float64), alpha.astype(np.float64), 0.1, rp, ci, vq.astype(np.float64), F, kernel, 3, 1.0 / F, coef0)
        assert nrm_rel(dec_o, dec) < 2e-5
        safe = np.abs(dec_o) > 1e-4 * np.abs(dec_o).max()
        assert np.array_equal(np.sign(dec[safe]), np.sign(dec_o[safe]))


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", KERNELS)
def test_csr_predict_and_update_w(kernel, dt):
    n, F, P = 600, 800, 1300
    rp, ci, v, y = make(n, F, 0.02, seed=31, dt=dt)
    prp, pci, pv, _ = make(P, F, 0.02, seed=32, dt=dt, min_nnz=0)
    alpha = (np.random.default_rng(7).standard_normal(n) * 0.1).astype(dt)
    with capi.Handle(dt, kernel, gamma=1.0 / F) as h:
        h.set_option(capi.OPT_SPARSE_BLOCK, 256)
        h.set_data_csr(rp, ci, v, F)
        dec = h.predict_csr(alpha, -0.2, prp, pci, pv)
        dec_o = port.csr_predict(rp, ci, v.astype(np.float64), alpha.astype(np.float64), -0.2, prp, pci, pv.astype(np.float64), F, kernel)
        assert nrm_rel(dec_o, dec) < (1e-12 if dt == np.float64 else 2e-5)
        if dt == np.float64:
            assert np.array_equal(h.predict_label(dec), port.sign(dec_o))  # bit-identical labels
        empty = h.predict_csr(alpha, -0.2, np.zeros(1, np.int64), np.zeros(0, np.int32), np.zeros(0, dt))
        assert empty.shape == (0,)
        if kernel == "linear":
            w = h.update_w(alpha)
            w_o = port.update_w(datagen.csr_to_dense(rp, ci, v, F).astype(np.float64), alpha.astype(np.float64))
            assert nrm_rel(w_o, w) < (1e-13 if dt == np.float64 else 1e-5)


@pytest.mark.parametrize("kernel", KERNELS)
def test_csr_predict_chunk_pipeline(kernel, monkeypatch):
    """The points are staged in chunks over two buffer sets (upload + norms of chunk i + 1 on a second stream while chunk i computes):
    many small chunks (test hook B200SVM_PREDICT_CHUNK = stored entries per chunk) must give exactly the one-chunk result."""
    n, F, P = 500, 600, 2100
    dt = np.float32
    rp, ci, v, y = make(n, F, 0.03, seed=41, dt=dt)
    prp, pci, pv, _ = make(P, F, 0.03, seed=42, dt=dt, min_nnz=0)
    alpha = (np.random.default_rng(8).standard_normal(n) * 0.1).astype(dt)
    with capi.Handle(dt, kernel, gamma=1.0 / F) as h:
        h.set_option(capi.OPT_SPARSE_BLOCK, 256)
        h.set_data_csr(rp, ci, v, F)
        one = h.predict_csr(alpha, 0.1, prp, pci, pv)
        for chunk in (1, 700, 5000):  # one row per chunk; ~40 rows; ~280 rows
            monkeypatch.setenv("B200SVM_PREDICT_CHUNK", str(chunk))
            many = h.predict_csr(alpha, 0.1, prp, pci, pv)
            # the fixed-point scale follows the largest norm of a CHUNK: the Gram sums round on a (slightly) different grid
            assert nrm_rel(one, many) < 2e-6, chunk
        monkeypatch.delenv("B200SVM_PREDICT_CHUNK")
        assert np.array_equal(one, h.predict_csr(alpha, 0.1, prp, pci, pv))


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", ["polynomial", "rbf"])
@pytest.mark.parametrize("shape", [(300, 2000, 0.1, 128, 1), (200, 1500, 0.4, 128, 1), (900, 500, 0.05, 128, 2), (1000, 3000, 0.01, 256, 1)])
def test_csr_long_rows_and_launch_groups(kernel, dt, shape):
    """Rows of hundreds of stored entries, dense fill of the accumulator and several launch groups (segment pointers restart per group)."""
    n, F, density, W, G = shape
    rp, ci, v, _ = make(n, F, density, seed=n, dt=dt)
    d = np.random.default_rng(n).uniform(-1, 2, n - 1).astype(dt)
    want = None
    for cache in (capi.CACHE_OFF, capi.CACHE_AUTO):
        with capi.Handle(dt, kernel, 2, 1.0 / F, 0.0, 1.0) as h:
            h.set_option(capi.OPT_SPARSE_BLOCK, W)
            h.set_option(capi.OPT_SPARSE_GROUP, G)
            h.set_option(capi.OPT_SPARSE_CACHE, cache)
            h.set_data_csr(rp, ci, v, F)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            h.set_q(q, qa)
            got = h.matvec(d, 1.0)
            again = h.matvec(d, 1.0)
            assert np.array_equal(got, again)  # bit-reproducible
        if want is None:
            want = port.csr_matvec(rp, ci, v.astype(np.float64), F, q.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 2, 1.0 / F, 0.0, 1.0)
        assert nrm_rel(want, got) < (1e-12 if dt == np.float64 else 2e-5), nrm_rel(want, got)


def test_fp22_launch_groups():
    n, F = 700, 3000
    rp, ci, v, _ = make(n, F, 0.01, seed=77, dt=np.float32)
    prp, packed = datagen.pack_fp22_csr(rp, ci, v)
    vq = datagen.fp22_round(v)
    d = np.random.default_rng(3).uniform(-1, 2, n - 1).astype(np.float32)
    outs = []
    for G in (1, 2, 0):
        with capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0) as h:
            h.set_option(capi.OPT_SPARSE_BLOCK, 128)
            h.set_option(capi.OPT_SPARSE_GROUP, G)
            h.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
            q = h.generate_q()
            h.set_q(q, h.kxx_last() + 1.0)
            outs.append(h.matvec(d, 1.0))
    want = port.csr_matvec(rp, ci, vq.astype(np.float64), F, q.astype(np.float64), d.astype(np.float64), float(h.np_dtype.type(2.0)), 1.0, "rbf", 3, 1.0 / F, 0.0, 1.0)
    for o in outs:
        assert nrm_rel(want, o) < 2e-5
    assert nrm_rel(outs[0], outs[2]) < 1e-6


def test_csr_row_too_long_is_rejected():
    n, F = 24, 12000
    rp, ci, v, _ = make(n, F, 0.5, seed=1, dt=np.float64)
    with capi.Handle(np.float64, "rbf", gamma=1.0 / F) as h:
        h.set_data_csr(rp, ci, v, F)
        q = h.generate_q()
        h.set_q(q, 2.0)
        with pytest.raises(capi.B200Error, match="4096 stored entries"):
            h.matvec(np.ones(n - 1))


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_csr_rows_of_thousands_of_entries(dt):
    """Rows beyond 4 x the default CTA size (the CTA is widened) and more than 1024 additions per accumulator slot."""
    n, F = 60, 5000
    rp, ci, v, _ = make(n, F, 0.5, seed=2, dt=dt)
    d = np.random.default_rng(2).uniform(-1, 2, n - 1).astype(dt)
    with capi.Handle(dt, "rbf", 3, 1.0 / F, 0.0, 1.0) as h:
        h.set_data_csr(rp, ci, v, F)
        q = h.generate_q()
        qa = h.kxx_last() + 1.0
        h.set_q(q, qa)
        got = h.matvec(d, 1.0)
    want = port.csr_matvec(rp, ci, v.astype(np.float64), F, q.astype(np.float64), d.astype(np.float64), qa, 1.0, "rbf", 3, 1.0 / F, 0.0, 1.0)
    assert nrm_rel(want, got) < (1e-12 if dt == np.float64 else 2e-5), nrm_rel(want, got)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
@pytest.mark.parametrize("kernel", ["polynomial", "rbf"])
def test_csr_dense_columns_take_the_long_segment_path(kernel, dt):
    """Two features present in (almost) every row: their segments exceed the group threshold and are walked by the whole CTA."""
    n, F = 1500, 4000
    rng = np.random.default_rng(11)
    X = np.zeros((n, F), dt)
    mask = rng.random((n, F)) < 0.004
    X[mask] = rng.uniform(-1, 1, int(mask.sum())).astype(dt)
    X[:, 7] = rng.uniform(-1, 1, n).astype(dt)
    X[::2, 1234] = rng.uniform(-1, 1, (n + 1) // 2).astype(dt)
    rp, ci, v = datagen.dense_to_csr(X)
    d = rng.uniform(-1, 2, n - 1).astype(dt)
    outs = []
    for lanes, threads in ((0, 0), (1, 128), (32, 1024), (4, 256)):
        with capi.Handle(dt, kernel, 2, 1.0 / 16, 0.5, 1.0) as h:
            h.set_option(capi.OPT_SPARSE_BLOCK, 512)
            h.set_option(capi.OPT_SPARSE_CACHE, capi.CACHE_OFF if lanes else capi.CACHE_AUTO)
            h.set_option(capi.OPT_SPARSE_LANES, lanes)
            h.set_option(capi.OPT_SPARSE_THREADS, threads)
            h.set_data_csr(rp, ci, v, F)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            h.set_q(q, qa)
            outs.append(h.matvec(d, 1.0))
    want = port.csr_matvec(rp, ci, v.astype(np.float64), F, q.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 2, 1.0 / 16, 0.5, 1.0)
    for o in outs:
        assert nrm_rel(want, o) < (1e-12 if dt == np.float64 else 2e-5), nrm_rel(want, o)


@pytest.mark.parametrize("dt", [np.float64, np.float32])
def test_csr_cache_one_pass_and_two_pass_assembly_agree(dt, monkeypatch):
    """The S cache is normally assembled in one pass (entries reserved per visit with an atomic cursor, capacity bounded by the
    column counts); B200SVM_CACHE_TWO_PASS=1 forces the exact count / prefix-sum / fill assembly.  Where a visit lands in memory
    differs, what it holds does not: the matvec is bit-identical."""
    n, F = 2500, 1500
    rp, ci, v, _ = make(n, F, 0.02, seed=21, dt=dt)
    d = np.random.default_rng(21).uniform(-1, 2, n - 1).astype(dt)
    outs, entries = [], []
    for two_pass in ("0", "1"):
        monkeypatch.setenv("B200SVM_CACHE_TWO_PASS", two_pass)
        with capi.Handle(dt, "rbf", gamma=1.0 / F) as h:
            h.set_option(capi.OPT_SPARSE_BLOCK, 256)
            h.set_option(capi.OPT_SPARSE_CACHE, capi.CACHE_REQUIRED)
            h.set_data_csr(rp, ci, v, F)
            q = h.generate_q()
            h.set_q(q, h.kxx_last() + 1.0)
            outs.append(h.matvec(d))
            entries.append(h.stats().sparse_cache_entries)
    assert entries[0] == entries[1] > 0
    assert np.array_equal(outs[0], outs[1])


def test_csr_cache_budget_falls_back_or_fails_loudly():
    n, F = 900, 600
    rp, ci, v, _ = make(n, F, 0.05, seed=9, dt=np.float32)
    d = np.random.default_rng(9).uniform(-1, 2, n - 1).astype(np.float32)
    outs = []
    for mode, mb in ((capi.CACHE_AUTO, 0), (capi.CACHE_AUTO, 1), (capi.CACHE_REQUIRED, 1)):
        with capi.Handle(np.float32, "rbf", gamma=1.0 / F) as h:
            h.set_option(capi.OPT_SPARSE_CACHE, mode)
            h.set_option(capi.OPT_SPARSE_CACHE_MB, mb)
            h.set_data_csr(rp, ci, v, F)
            q = h.generate_q()
            h.set_q(q, 2.0)
            if mode == capi.CACHE_REQUIRED:
                with pytest.raises(capi.B200Error, match="do not fit"):
                    h.matvec(d)
                continue
            outs.append(h.matvec(d))
            assert (h.stats().sparse_cache_entries > 0) == (mb == 0)  # a 1 MiB budget cannot hold the fixed 64 MiB reserve
    assert nrm_rel(outs[0], outs[1]) < 2e-6


def test_csr_predict_rejects_bad_points():
    """Column indices of predict points are validated on the device together with the norms (no host pass over the entries)."""
    n, F = 200, 300
    rp, ci, v, _ = make(n, F, 0.05, seed=4, dt=np.float64)
    alpha = np.ones(n)
    with capi.Handle(np.float64, "rbf", gamma=1.0 / F) as h:
        h.set_data_csr(rp, ci, v, F)
        prp = np.array([0, 2, 3], np.int64)
        ok = h.predict_csr(alpha, 0.0, prp, np.array([1, 5, 0], np.int32), np.ones(3))
        assert ok.shape == (2,)
        for bad_cols in ([1, F, 0], [5, 1, 0], [-1, 2, 0]):  # out of range, not ascending, negative
            with pytest.raises(capi.B200Error, match="Number of features per data point must match"):
                h.predict_csr(alpha, 0.0, prp, np.array(bad_cols, np.int32), np.ones(3))
        # float points into a double handle are converted on the host
        dec32 = h.predict_csr(alpha, 0.0, prp, np.array([1, 5, 0], np.int32), np.ones(3, np.float32))
        assert np.array_equal(dec32, ok)


def test_csr_errors_are_loud():
    with capi.Handle(np.float64, "rbf", gamma=0.5) as h:
        rp = np.array([0, 2, 3], np.int64)
        with pytest.raises(capi.B200Error, match="sorted and unique"):
            h.set_data_csr(rp, np.array([1, 1, 0], np.int32), np.ones(3), 4)
        with pytest.raises(capi.B200Error, match="out of range"):
            h.set_data_csr(rp, np.array([0, 9, 0], np.int32), np.ones(3), 4)
        with pytest.raises(capi.B200Error, match="float handle"):
            h.set_data_csr(rp, None, np.ones(3, np.uint32), 4, value_format=capi.VAL_FP22)


@pytest.mark.parametrize("kernel", ["polynomial", "rbf"])
@pytest.mark.parametrize("shape", [(1100, 300, 0.05, 256), (2500, 200, 0.08, 1024), (3000, 150, 0.1, 0), (130, 60, 0.2, 256)])
@pytest.mark.parametrize("fp22", [False, True])
def test_csr_dense_tile_store_of_a_filled_S(kernel, shape, fp22):
    """Float, S densely filled: S is kept as dense lower-triangle 128 x 256 tiles (B200SVM_OPT_SPARSE_TILES) and swept like the
    dense kernel-matrix cache.  Against the oracle's CSR restatement and against the indexed cache on the same handle data."""
    n, F, density, W = shape
    rp, ci, v, y = make(n, F, density, seed=n + F, dt=np.float32, min_nnz=0)
    fmt = None
    if fp22:
        v = datagen.fp22_round(v)
        rp_s, v_s = datagen.pack_fp22_csr(rp, ci, v)
        ci_s, fmt = None, capi.VAL_FP22
    else:
        rp_s, ci_s, v_s = rp, ci, v
    coef0 = 0.5 if kernel == "polynomial" else 0.0
    rng = np.random.default_rng(n)
    d = rng.uniform(-1, 2, n - 1).astype(np.float32)
    res = {}
    for tiles in (2, 1):
        with capi.Handle(np.float32, kernel, 3, 1.0 / F, coef0, 1.5) as h:
            if W:
                h.set_option(capi.OPT_SPARSE_BLOCK, W)
            h.set_option(capi.OPT_SPARSE_CACHE, capi.CACHE_REQUIRED)
            h.set_option(capi.OPT_SPARSE_TILES, tiles)
            h.set_data_csr(rp_s, ci_s, v_s, F, value_format=fmt)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0 / 1.5
            h.set_q(q, qa)
            got = h.matvec(d, 1.0)
            got_b = h.matvec(d, 1.0)   # second matvec: the sweep over the stored tiles / entries
            st = h.stats()
            assert (st.dense_cache_bytes > 0) == (tiles == 2), (tiles, st.dense_cache_bytes)
            assert np.array_equal(got, got_b)
            alpha, rho, it = h.learn(y, eps=1e-3)
            res[tiles] = (q, qa, got, alpha, it)
    q, qa, got, alpha, it = res[2]
    want = port.csr_matvec(rp, ci, v.astype(np.float64), F, q.astype(np.float64), d.astype(np.float64), qa, 1.0, kernel, 3, 1.0 / F, coef0, 1.5)
    assert nrm_rel(want, got) < 2e-5, nrm_rel(want, got)
    assert nrm_rel(res[1][2], got) < 2e-5
    assert it == res[1][4]
    assert nrm_rel(res[1][3], alpha) < 1e-3
