"""TEST INFRASTRUCTURE — ctypes binding of oracle/lssvm_oracle.c (plain-C restatement).  NOT product code."""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "liblssvm_oracle.so")
_LIB = None

KERNELS = {"linear": 0, "polynomial": 1, "rbf": 2}


def build(force: bool = False) -> str:
    src = os.path.join(_HERE, "lssvm_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "port"], stdout=subprocess.DEVNULL)
    return _SO


def lib():
    global _LIB
    if _LIB is None:
        if not os.path.exists(_SO):
            build()
        L = C.CDLL(_SO)
        vp, i, l = C.c_void_p, C.c_int, C.c_long
        for sfx, r in (("_f32", C.c_float), ("_f64", C.c_double)):
            f = getattr(L, "lssvm_kernel" + sfx)
            f.restype = r
            f.argtypes = [i, vp, vp, l, i, r, r]
            getattr(L, "lssvm_generate_q" + sfx).argtypes = [i, vp, l, l, i, r, r, vp]
            getattr(L, "lssvm_matvec" + sfx).argtypes = [i, vp, l, l, i, r, r, vp, vp, r, r, r, vp]
            getattr(L, "lssvm_matvec_rows" + sfx).argtypes = [i, vp, l, l, i, r, r, vp, vp, r, r, r, l, l, vp]
            f = getattr(L, "lssvm_solver_cg" + sfx)
            f.restype = l
            f.argtypes = [i, vp, l, l, i, r, r, r, vp, vp, r, l, r, i, vp, vp]
            f = getattr(L, "lssvm_learn" + sfx)
            f.restype = l
            f.argtypes = [i, vp, vp, l, l, i, r, r, r, r, i, vp, vp]
            getattr(L, "lssvm_update_w" + sfx).argtypes = [vp, vp, l, l, vp]
            getattr(L, "lssvm_predict" + sfx).argtypes = [i, vp, vp, l, l, i, r, r, r, vp, l, vp]
            getattr(L, "lssvm_sign" + sfx).argtypes = [vp, l, vp]
            getattr(L, "lssvm_csr_generate_q" + sfx).argtypes = [i, vp, vp, vp, l, i, r, r, vp]
            getattr(L, "lssvm_csr_matvec_rows" + sfx).argtypes = [i, vp, vp, vp, l, i, r, r, vp, vp, r, r, r, l, l, vp]
            getattr(L, "lssvm_csr_predict" + sfx).argtypes = [i, vp, vp, vp, vp, l, i, r, r, r, vp, vp, vp, l, vp]
        L.lssvm_fp22_encode.restype = C.c_uint32
        L.lssvm_fp22_encode.argtypes = [C.c_float]
        L.lssvm_fp22_decode.restype = C.c_float
        L.lssvm_fp22_decode.argtypes = [C.c_uint32]
        L.lssvm_fp22_round.restype = C.c_float
        L.lssvm_fp22_round.argtypes = [C.c_float]
        _LIB = L
    return _LIB


def _sfx(a):
    return {np.dtype(np.float32): "_f32", np.dtype(np.float64): "_f64"}[a.dtype]


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def _g(F, gamma):
    return 1.0 / F if gamma is None else gamma


def kernel_function(a, b, kernel, degree=3, gamma=1.0, coef0=0.0):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b, a.dtype)
    return getattr(lib(), "lssvm_kernel" + _sfx(a))(KERNELS[kernel], _p(a), _p(b), a.shape[0], degree, gamma, coef0)


def generate_q(X, kernel, degree=3, gamma=None, coef0=0.0):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    q = np.empty(n - 1, X.dtype)
    getattr(lib(), "lssvm_generate_q" + _sfx(X))(KERNELS[kernel], _p(X), n, F, degree, _g(F, gamma), coef0, _p(q))
    return q


def matvec(X, q, d, QA_cost, add, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, ret=None, rows=None):
    """ret += add * Abar * d.  rows=None: the reference's serial mirrored loop; rows=(lo, hi): parallel full-row form."""
    X = np.ascontiguousarray(X)
    n, F = X.shape
    q = np.ascontiguousarray(q, X.dtype)
    d = np.ascontiguousarray(d, X.dtype)
    out = np.zeros(n - 1, X.dtype) if ret is None else np.array(ret, X.dtype, copy=True)
    if rows is None:
        getattr(lib(), "lssvm_matvec" + _sfx(X))(KERNELS[kernel], _p(X), n, F, degree, _g(F, gamma), coef0, _p(q), _p(d), QA_cost, 1.0 / cost, add, _p(out))
    else:
        getattr(lib(), "lssvm_matvec_rows" + _sfx(X))(KERNELS[kernel], _p(X), n, F, degree, _g(F, gamma), coef0, _p(q), _p(d), QA_cost, 1.0 / cost, add, rows[0], rows[1], _p(out))
    return out


def solver_cg(X, b, q, QA_cost, imax, eps, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, use_rows=False):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    b = np.ascontiguousarray(b, X.dtype)
    q = np.ascontiguousarray(q, X.dtype)
    alpha = np.empty(n - 1, X.dtype)
    delta = np.zeros(1, X.dtype)
    it = getattr(lib(), "lssvm_solver_cg" + _sfx(X))(KERNELS[kernel], _p(X), n, F, degree, _g(F, gamma), coef0, cost, _p(b), _p(q), QA_cost, imax, eps, int(use_rows), _p(alpha), _p(delta))
    return alpha, it, float(delta[0])


def learn(X, y, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, eps=1e-3, use_rows=False):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    y = np.ascontiguousarray(y, X.dtype)
    alpha = np.empty(n, X.dtype)
    rho = np.zeros(1, X.dtype)
    it = getattr(lib(), "lssvm_learn" + _sfx(X))(KERNELS[kernel], _p(X), _p(y), n, F, degree, _g(F, gamma), coef0, cost, eps, int(use_rows), _p(alpha), _p(rho))
    return alpha, float(rho[0]), it


def update_w(X, alpha):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    alpha = np.ascontiguousarray(alpha, X.dtype)
    w = np.empty(F, X.dtype)
    getattr(lib(), "lssvm_update_w" + _sfx(X))(_p(X), _p(alpha), n, F, _p(w))
    return w


def predict(X, alpha, rho, points, kernel, degree=3, gamma=None, coef0=0.0):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    alpha = np.ascontiguousarray(alpha, X.dtype)
    points = np.ascontiguousarray(points, X.dtype)
    out = np.empty(points.shape[0], X.dtype)
    getattr(lib(), "lssvm_predict" + _sfx(X))(KERNELS[kernel], _p(X), _p(alpha), n, F, degree, _g(F, gamma), coef0, rho, _p(points), points.shape[0], _p(out))
    return out


def sign(v):
    v = np.ascontiguousarray(v)
    out = np.empty_like(v)
    getattr(lib(), "lssvm_sign" + _sfx(v))(_p(v), v.shape[0], _p(out))
    return out


# ---- CSR ----
def _csr(rp, ci, v):
    return np.ascontiguousarray(rp, np.int64), np.ascontiguousarray(ci, np.int32), np.ascontiguousarray(v)


def csr_generate_q(rp, ci, v, F, kernel, degree=3, gamma=None, coef0=0.0):
    rp, ci, v = _csr(rp, ci, v)
    n = rp.shape[0] - 1
    q = np.empty(n - 1, v.dtype)
    getattr(lib(), "lssvm_csr_generate_q" + _sfx(v))(KERNELS[kernel], _p(rp), _p(ci), _p(v), n, degree, _g(F, gamma), coef0, _p(q))
    return q


def csr_matvec(rp, ci, v, F, q, d, QA_cost, add, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, ret=None, rows=None):
    rp, ci, v = _csr(rp, ci, v)
    n = rp.shape[0] - 1
    q = np.ascontiguousarray(q, v.dtype)
    d = np.ascontiguousarray(d, v.dtype)
    out = np.zeros(n - 1, v.dtype) if ret is None else np.array(ret, v.dtype, copy=True)
    lo, hi = (0, n - 1) if rows is None else rows
    getattr(lib(), "lssvm_csr_matvec_rows" + _sfx(v))(KERNELS[kernel], _p(rp), _p(ci), _p(v), n, degree, _g(F, gamma), coef0, _p(q), _p(d), QA_cost, 1.0 / cost, add, lo, hi, _p(out))
    return out


def csr_predict(rp, ci, v, alpha, rho, prp, pci, pv, F, kernel, degree=3, gamma=None, coef0=0.0):
    rp, ci, v = _csr(rp, ci, v)
    prp, pci, pv = _csr(prp, pci, pv)
    pv = pv.astype(v.dtype, copy=False)
    n = rp.shape[0] - 1
    npts = prp.shape[0] - 1
    alpha = np.ascontiguousarray(alpha, v.dtype)
    out = np.empty(npts, v.dtype)
    getattr(lib(), "lssvm_csr_predict" + _sfx(v))(KERNELS[kernel], _p(rp), _p(ci), _p(v), _p(alpha), n, degree, _g(F, gamma), coef0, rho, _p(prp), _p(pci), _p(pv), npts, _p(out))
    return out


# ---- FP22 ----
def fp22_encode(x: float) -> int:
    return int(lib().lssvm_fp22_encode(float(np.float32(x))))


def fp22_decode(v: int) -> float:
    return float(lib().lssvm_fp22_decode(int(v)))


def fp22_round(x: float) -> float:
    return float(lib().lssvm_fp22_round(float(np.float32(x))))
