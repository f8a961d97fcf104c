"""ctypes binding of the C ABI (include/b200svm.h).  Test/bench harness only — the product is libb200svm.so itself.

There is no CPU fallback: loading fails loudly when the library has not been built, and every compute call fails with
the library's own error text when no sm_100 device is usable.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libb200svm.so")

F32, F64 = 0, 1
KERNELS = {"linear": 0, "polynomial": 1, "rbf": 2}
VAL_F32, VAL_F64, VAL_FP22 = 0, 1, 2
OPT_MATVEC_PATH, OPT_SYNC_EVERY, OPT_SPARSE_BLOCK, OPT_SPARSE_GROUP, OPT_SPARSE_LANES, OPT_TC_MODE, OPT_SPARSE_THREADS = 1, 2, 3, 4, 5, 6, 7
OPT_SPARSE_CACHE, OPT_SPARSE_CACHE_MB, OPT_DENSE_CACHE, OPT_PEER_EXCHANGE = 8, 9, 10, 11
OPT_SPARSE_TILES = 12
NOALLOC_SCRATCH_BYTES = 4 << 20  # B200SVM_NOALLOC_SCRATCH_BYTES
CACHE_AUTO, CACHE_OFF, CACHE_REQUIRED = 0, 1, 2
TC_F16X3, TC_FP8X = 1, 2
PATH_AUTO, PATH_SIMT, PATH_TC = 0, 1, 2

# every symbol include/b200svm.h declares
SYMBOLS = [
    "b200svm_last_error", "b200svm_version", "b200svm_create", "b200svm_destroy", "b200svm_set_stream",
    "b200svm_set_data_dense", "b200svm_set_data_csr", "b200svm_generate_q", "b200svm_get_kxx_last", "b200svm_set_q",
    "b200svm_matvec", "b200svm_cg_solve", "b200svm_learn", "b200svm_update_w", "b200svm_predict_dense",
    "b200svm_predict_csr", "b200svm_comm_unique_id", "b200svm_comm_init", "b200svm_set_option", "b200svm_get_stats",
    "b200svm_reset_stats", "b200svm_partition_rows", "b200svm_partition_tiles", "b200svm_probe_fma_tflops", "b200svm_probe_smem_atomic_gops", "b200svm_partition_tile_rows", "b200svm_canonical_groups",
    "b200svm_update_w_noalloc",
]


class Stats(C.Structure):
    _fields_ = [
        ("kernel_launches", C.c_int64),
        ("matvecs", C.c_int64),
        ("matvec_ms", C.c_double),
        ("matvec_path", C.c_int64),
        ("h2d_bytes", C.c_int64),
        ("d2h_bytes", C.c_int64),
        ("collectives", C.c_int64),
        ("collective_bytes", C.c_int64),
        ("peer_exchanges", C.c_int64),
        ("matvecs_timed", C.c_int64),
        ("cache_build_ms", C.c_double),
        ("dense_cache_bytes", C.c_int64),
        ("sparse_cache_entries", C.c_int64),
        ("cg_vector_ms", C.c_double),
        ("cg_vector_timed", C.c_int64),
    ]


class B200Error(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[b200svm error {code}] {msg}")
        self.code = code
        self.message = msg


_LIB = None


def load():
    global _LIB
    if _LIB is not None:
        return _LIB
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(f"{LIB_PATH} is missing — build it with `python -m plssvm_sparse_fp22_b200.build` "
                           "(__graft_entry__.build()); there is no CPU fallback")
    L = C.CDLL(LIB_PATH)
    vp, i, d, i64 = C.c_void_p, C.c_int, C.c_double, C.c_int64
    pi64, pd = C.POINTER(i64), C.POINTER(d)
    L.b200svm_last_error.restype = C.c_char_p
    L.b200svm_last_error.argtypes = [vp]
    L.b200svm_version.restype = i
    L.b200svm_create.argtypes = [C.POINTER(vp), i, i, i, d, d, d, i]
    L.b200svm_destroy.argtypes = [vp]
    L.b200svm_set_stream.argtypes = [vp, vp]
    L.b200svm_set_data_dense.argtypes = [vp, vp, i64, i64, i64]
    L.b200svm_set_data_csr.argtypes = [vp, vp, vp, vp, i64, i64, i64, i]
    L.b200svm_generate_q.argtypes = [vp, vp]
    L.b200svm_get_kxx_last.argtypes = [vp, pd]
    L.b200svm_set_q.argtypes = [vp, vp, d]
    L.b200svm_matvec.argtypes = [vp, vp, d, vp]
    L.b200svm_cg_solve.argtypes = [vp, vp, vp, d, i64, d, vp, pi64, pd]
    L.b200svm_learn.argtypes = [vp, vp, d, vp, pd, pi64]
    L.b200svm_update_w.argtypes = [vp, vp, vp]
    L.b200svm_predict_dense.argtypes = [vp, vp, d, vp, i64, i64, vp]
    L.b200svm_predict_csr.argtypes = [vp, vp, d, vp, vp, vp, i64, i64, i, vp]
    L.b200svm_comm_unique_id.argtypes = [vp]
    L.b200svm_comm_init.argtypes = [vp, vp, i, i]
    L.b200svm_set_option.argtypes = [vp, i, i64]
    L.b200svm_get_stats.argtypes = [vp, C.POINTER(Stats)]
    L.b200svm_reset_stats.argtypes = [vp]
    L.b200svm_probe_fma_tflops.argtypes = [vp, i, pd]
    L.b200svm_probe_smem_atomic_gops.argtypes = [vp, pd]
    L.b200svm_partition_rows.argtypes = [i64, i, i, pi64, pi64]
    L.b200svm_partition_tiles.argtypes = [i64, i, i, pi64, pi64, pi64]
    L.b200svm_partition_tile_rows.argtypes = [i64, i, i, pi64, pi64]
    L.b200svm_canonical_groups.argtypes = [i]
    L.b200svm_update_w_noalloc.argtypes = [i, vp, i64, i64, i64, vp, vp, vp, C.c_char_p, C.c_size_t]
    _LIB = L
    return L


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Handle:
    """Thin owner of a b200svm_handle.  Method names mirror the five backend virtuals of plssvm::csvm<T>
    (include/plssvm/csvm.hpp:188-214): setup_data_on_device, generate_q, solver_CG, update_w, predict."""

    def __init__(self, dtype, kernel="linear", degree=3, gamma=None, coef0=0.0, cost=1.0, device=0):
        self.L = load()
        self.np_dtype = np.dtype(dtype)
        self.dtype = {np.dtype(np.float32): F32, np.dtype(np.float64): F64}[self.np_dtype]
        self.kernel = kernel
        self.gamma = gamma
        self.h = C.c_void_p()
        self._args = (KERNELS[kernel], degree, coef0, cost, device)
        self._created = False
        if gamma is not None:
            self._create(gamma)

    def _create(self, gamma):
        k, degree, coef0, cost, device = self._args
        rc = self.L.b200svm_create(C.byref(self.h), self.dtype, k, degree, float(gamma), coef0, cost, device)
        if rc != 0:
            raise B200Error(rc, self.L.b200svm_last_error(None).decode())
        self.gamma = gamma
        self._created = True

    def _chk(self, rc):
        if rc != 0:
            raise B200Error(rc, self.L.b200svm_last_error(self.h if self._created else None).decode())

    def close(self):
        if self._created:
            self.L.b200svm_destroy(self.h)
            self._created = False

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # -- setup_data_on_device ------------------------------------------------------------------------------------
    def set_data(self, X):
        X = np.ascontiguousarray(X, self.np_dtype)
        if not self._created:
            self._create(1.0 / X.shape[1])  # default gamma = 1 / num_features (src/plssvm/parameter.cpp:150-152)
        self.n, self.F = X.shape
        self._chk(self.L.b200svm_set_data_dense(self.h, _p(X), self.n, self.F, self.F))

    def set_data_csr(self, rowptr, colidx, values, F, value_format=None):
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        n = rowptr.shape[0] - 1
        if not self._created:
            self._create(1.0 / F)
        if value_format is None:
            value_format = VAL_F32 if values.dtype == np.float32 else VAL_F64
        if value_format == VAL_FP22:
            values = np.ascontiguousarray(values, np.uint32)
            ci = None
        else:
            values = np.ascontiguousarray(values)
            ci = np.ascontiguousarray(colidx, np.int32)
        self.n, self.F = n, F
        self._chk(self.L.b200svm_set_data_csr(self.h, _p(rowptr), _p(ci), _p(values), n, F, int(rowptr[-1]), value_format))

    def set_stream(self, stream_ptr):
        self._chk(self.L.b200svm_set_stream(self.h, C.c_void_p(stream_ptr)))

    def set_option(self, opt, value):
        if not self._created:
            raise B200Error(1, "handle not created yet: pass gamma to Handle(...) or call set_data first")
        self._chk(self.L.b200svm_set_option(self.h, opt, int(value)))

    # -- generate_q ----------------------------------------------------------------------------------------------
    def generate_q(self):
        q = np.empty(self.n - 1, self.np_dtype)
        self._chk(self.L.b200svm_generate_q(self.h, _p(q)))
        return q

    def kxx_last(self):
        v = C.c_double(0)
        self._chk(self.L.b200svm_get_kxx_last(self.h, C.byref(v)))
        return v.value

    def set_q(self, q, QA_cost):
        q = None if q is None else np.ascontiguousarray(q, self.np_dtype)
        self._chk(self.L.b200svm_set_q(self.h, _p(q), float(QA_cost)))

    # -- run_device_kernel ---------------------------------------------------------------------------------------
    def matvec(self, d, add=1.0, ret=None):
        d = np.ascontiguousarray(d, self.np_dtype)
        out = np.zeros(self.n - 1, self.np_dtype) if ret is None else np.array(ret, self.np_dtype, copy=True)
        self._chk(self.L.b200svm_matvec(self.h, _p(d), float(add), _p(out)))
        return out

    # -- solver_CG -----------------------------------------------------------------------------------------------
    def solver_cg(self, b, q, QA_cost, imax, eps):
        b = np.ascontiguousarray(b, self.np_dtype)
        q = None if q is None else np.ascontiguousarray(q, self.np_dtype)
        alpha = np.empty(self.n - 1, self.np_dtype)
        it = C.c_int64(0)
        delta = C.c_double(0)
        self._chk(self.L.b200svm_cg_solve(self.h, _p(b), _p(q), float(QA_cost), int(imax), float(eps), _p(alpha), C.byref(it), C.byref(delta)))
        return alpha, it.value, delta.value

    def learn(self, y, eps=1e-3):
        y = np.ascontiguousarray(y, self.np_dtype)
        alpha = np.empty(self.n, self.np_dtype)
        rho = C.c_double(0)
        it = C.c_int64(0)
        self._chk(self.L.b200svm_learn(self.h, _p(y), float(eps), _p(alpha), C.byref(rho), C.byref(it)))
        return alpha, rho.value, it.value

    # -- update_w / predict ----------------------------------------------------------------------------------------
    def update_w(self, alpha):
        alpha = np.ascontiguousarray(alpha, self.np_dtype)
        w = np.empty(self.F, self.np_dtype)
        self._chk(self.L.b200svm_update_w(self.h, _p(alpha), _p(w)))
        return w

    def predict(self, alpha, rho, points):
        alpha = np.ascontiguousarray(alpha, self.np_dtype)
        points = np.ascontiguousarray(points, self.np_dtype)
        out = np.empty(points.shape[0], self.np_dtype)
        ld = points.shape[1] if points.ndim == 2 else self.F
        self._chk(self.L.b200svm_predict_dense(self.h, _p(alpha), float(rho), _p(points), points.shape[0], ld, _p(out)))
        return out

    def predict_csr(self, alpha, rho, rowptr, colidx, values, value_format=None):
        alpha = np.ascontiguousarray(alpha, self.np_dtype)
        rowptr = np.ascontiguousarray(rowptr, np.int64)
        P = rowptr.shape[0] - 1
        if value_format is None:
            value_format = VAL_F32 if values.dtype == np.float32 else VAL_F64
        if value_format == VAL_FP22:
            values = np.ascontiguousarray(values, np.uint32)
            ci = None
        else:
            values = np.ascontiguousarray(values)
            ci = np.ascontiguousarray(colidx, np.int32)
        out = np.empty(P, self.np_dtype)
        self._chk(self.L.b200svm_predict_csr(self.h, _p(alpha), float(rho), _p(rowptr), _p(ci), _p(values), P, int(rowptr[-1]), value_format, _p(out)))
        return out

    @staticmethod
    def predict_label(decision):
        """operators::sign (include/plssvm/detail/operators.hpp:174-177): x > 0 ? +1 : -1."""
        return np.where(decision > 0, 1.0, -1.0).astype(decision.dtype)

    # -- multi-GPU -------------------------------------------------------------------------------------------------
    def comm_init(self, id128: bytes, rank: int, world: int):
        buf = C.create_string_buffer(id128, 128)
        self._chk(self.L.b200svm_comm_init(self.h, buf, rank, world))

    # -- stats -----------------------------------------------------------------------------------------------------
    def stats(self) -> Stats:
        s = Stats()
        self._chk(self.L.b200svm_get_stats(self.h, C.byref(s)))
        return s

    def reset_stats(self):
        self._chk(self.L.b200svm_reset_stats(self.h))

    def probe_smem_atomic_gops(self):
        """Measured rate (1e9 / s) of 32-bit shared-memory atomic adds to spread addresses on the handle's device."""
        v = C.c_double(0)
        self._chk(self.L.b200svm_probe_smem_atomic_gops(self.h, C.byref(v)))
        return v.value

    def probe_fma_tflops(self, dtype=None):
        """Measured FMA peak (TFLOP/s) of the handle's device for float32 / float64."""
        v = C.c_double(0)
        dt = self.dtype if dtype is None else {np.dtype(np.float32): F32, np.dtype(np.float64): F64}[np.dtype(dtype)]
        self._chk(self.L.b200svm_probe_fma_tflops(self.h, dt, C.byref(v)))
        return v.value


def partition_rows(m: int, rank: int, world: int):
    lo, hi = C.c_int64(0), C.c_int64(0)
    rc = load().b200svm_partition_rows(m, rank, world, C.byref(lo), C.byref(hi))
    if rc != 0:
        raise B200Error(rc, "invalid partition request")
    return lo.value, hi.value


def partition_tiles(n: int, rank: int, world: int):
    first, count, total = C.c_int64(0), C.c_int64(0), C.c_int64(0)
    rc = load().b200svm_partition_tiles(n, rank, world, C.byref(first), C.byref(count), C.byref(total))
    if rc != 0:
        raise B200Error(rc, "invalid partition request")
    return first.value, count.value, total.value


def partition_tile_rows(m: int, rank: int, world: int):
    lo, hi = C.c_int64(0), C.c_int64(0)
    rc = load().b200svm_partition_tile_rows(m, rank, world, C.byref(lo), C.byref(hi))
    if rc != 0:
        raise B200Error(rc, "invalid partition request")
    return lo.value, hi.value


def canonical_groups(world: int) -> int:
    return load().b200svm_canonical_groups(world)


def comm_unique_id() -> bytes:
    L = load()
    buf = C.create_string_buffer(128)
    rc = L.b200svm_comm_unique_id(buf)
    if rc != 0:
        raise B200Error(rc, "ncclGetUniqueId failed")
    return buf.raw
