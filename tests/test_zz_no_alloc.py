"""b200svm_update_w_noalloc (include/b200svm.h): the allocation-free variant of update_w whose chunk partials live in a file-scope static
__device__ array.  No counterpart in the reference; kept in a file of its own that
pytest collects LAST so that a failure here cannot hide another test under `pytest -x`.

Checked through the C ABI against the existing entry point b200svm_update_w on the same data, with the tolerance the update_w parity
test uses (tests/test_gpu_parity.py: norm-wise 1e-13 double / 1e-5 float), on several calls in a row that share the static array
(different shapes and dtypes, so a stale or mis-addressed scratch would show)."""
import ctypes as C

import numpy as np
import pytest

from conftest import nrm_rel
from plssvm_sparse_fp22_b200 import capi, datagen

pytestmark = pytest.mark.gpu
TOL = {np.float64: 1e-13, np.float32: 1e-5}


def reference_w(X, alpha):
    """w from the existing entry point (handle + b200svm_update_w, host buffers)."""
    with capi.Handle(X.dtype.type, "linear") as h:
        h.set_data(X)
        return h.update_w(alpha)


def noalloc_w(torch, X, alpha, pad, stream=None):
    """w from b200svm_update_w_noalloc on device tensors allocated HERE (the entry point itself allocates nothing).  The rows are
    stored with leading dimension F + pad (padding poisoned with NaN), w_out carries one guard element behind its F reals."""
    L = capi.load()
    n, F = X.shape
    ld = F + pad
    Xd = torch.full((n, ld), float("nan"), dtype=torch.from_numpy(X).dtype, device="cuda")
    Xd[:, :F] = torch.from_numpy(X).cuda()
    ad = torch.from_numpy(alpha).cuda()
    wd = torch.full((F + 1,), -7.0, dtype=Xd.dtype, device="cuda")
    torch.cuda.synchronize()
    err = C.create_string_buffer(512)
    rc = L.b200svm_update_w_noalloc(capi.F64 if X.dtype == np.float64 else capi.F32, Xd.data_ptr(), n, F, ld, ad.data_ptr(),
                                    wd.data_ptr(), stream, err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    out = wd.cpu().numpy()
    assert out[F] == -7.0  # nothing written behind the F reals
    return out[:F]


def test_update_w_noalloc_matches_update_w_on_calls_in_a_row():
    import torch
    cases = [(np.float32, 3000, 200, 8, 1), (np.float32, 700, 50, 0, 2), (np.float64, 1500, 333, 3, 3), (np.float32, 3000, 200, 8, 1),
             (np.float64, 64, 5, 0, 4)]
    first = {}
    free_after = []
    for dt, n, F, pad, seed in cases:
        X, _ = datagen.make_dense(n, F, seed=seed, problem="planes", dtype=dt)
        alpha = np.random.default_rng(seed).standard_normal(n).astype(dt)
        w_ref = reference_w(X, alpha)
        w = noalloc_w(torch, X, alpha, pad)
        assert np.all(np.isfinite(w))
        assert nrm_rel(w_ref, w) < TOL[dt], (dt, n, F)
        key = (dt, n, F, pad, seed)
        if key in first:  # the same call again after other shapes used the array: bit-identical
            assert np.array_equal(first[key], w)
        first[key] = w
        # two more calls on identical inputs: the shared array keeps no state that leaks into the result, and the device's free
        # memory does not move between them (no hidden cudaMalloc)
        w2 = noalloc_w(torch, X, alpha, pad)
        free_after.append(torch.cuda.mem_get_info()[0])
        w3 = noalloc_w(torch, X, alpha, pad)
        free_after.append(torch.cuda.mem_get_info()[0])
        assert np.array_equal(w, w2) and np.array_equal(w, w3)
        assert abs(free_after[-1] - free_after[-2]) < (1 << 20)


def test_update_w_noalloc_on_a_side_stream():
    import torch
    X, _ = datagen.make_dense(2048, 96, seed=9, problem="planes", dtype=np.float32)
    alpha = np.random.default_rng(9).standard_normal(2048).astype(np.float32)
    s = torch.cuda.Stream()
    torch.cuda.synchronize()
    w = noalloc_w(torch, X, alpha, 0, stream=s.cuda_stream)
    assert nrm_rel(reference_w(X, alpha), w) < TOL[np.float32]


def test_update_w_noalloc_rejects_what_it_cannot_do():
    import torch
    L = capi.load()
    err = C.create_string_buffer(512)
    ld = capi.NOALLOC_SCRATCH_BYTES // 4 + 1  # one float row of partials no longer fits the static array
    Xd = torch.zeros((1, ld), dtype=torch.float32, device="cuda")
    ad = torch.ones(1, dtype=torch.float32, device="cuda")
    wd = torch.zeros(ld, dtype=torch.float32, device="cuda")
    rc = L.b200svm_update_w_noalloc(capi.F32, Xd.data_ptr(), 1, ld, ld, ad.data_ptr(), wd.data_ptr(), None, err, len(err))
    assert rc == 1 and b"exceed the static array" in err.value
    # the largest leading dimension that does fit still works
    ld -= 1
    rc = L.b200svm_update_w_noalloc(capi.F32, Xd.data_ptr(), 1, ld, ld, ad.data_ptr(), wd.data_ptr(), None, err, len(err))
    assert rc == 0, err.value.decode()
    torch.cuda.synchronize()
    # host pointers, null pointers, bad sizes / dtype: clear errors, no crash
    host = np.zeros(16, np.float32)
    hp = host.ctypes.data_as(C.c_void_p)
    assert L.b200svm_update_w_noalloc(capi.F32, hp, 4, 4, 4, ad.data_ptr(), wd.data_ptr(), None, err, len(err)) == 1
    assert b"device pointers" in err.value
    assert L.b200svm_update_w_noalloc(capi.F32, None, 4, 4, 4, ad.data_ptr(), wd.data_ptr(), None, err, len(err)) == 1
    assert L.b200svm_update_w_noalloc(capi.F32, Xd.data_ptr(), 1, 8, 4, ad.data_ptr(), wd.data_ptr(), None, err, len(err)) == 1
    assert L.b200svm_update_w_noalloc(7, Xd.data_ptr(), 1, 4, 4, ad.data_ptr(), wd.data_ptr(), None, None, 0) == 1
