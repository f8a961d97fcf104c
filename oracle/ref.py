"""TEST INFRASTRUCTURE — ctypes binding of oracle/_ref (the unmodified reference OpenMP backend).  NOT product code.

The shared objects are built by ``make -C oracle ref`` from the sources under /root/reference (see oracle/Makefile
and oracle/ref_driver.cpp).  Two ISA levels exist (x86-64-v3 / -v4); the one the host CPU supports is loaded.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None

KERNELS = {"linear": 0, "polynomial": 1, "rbf": 2}


def _cpu_has_avx512() -> bool:
    try:
        with open("/proc/cpuinfo") as f:
            txt = f.read()
        flags = txt.split("flags", 1)[1].split("\n", 1)[0]
        return all(k in flags for k in ("avx512f", "avx512bw", "avx512cd", "avx512dq", "avx512vl"))
    except Exception:
        return False


def lib_path() -> str:
    level = "v4" if _cpu_has_avx512() else "v3"
    return os.path.join(_HERE, "_ref", f"liboracle_ref_{level}.so")


def available() -> bool:
    return os.path.exists(lib_path())


def lib():
    global _LIB
    if _LIB is None:
        path = lib_path()
        if not os.path.exists(path):
            raise RuntimeError(f"{path} missing: run `make -C oracle ref` where /root/reference is present")
        L = C.CDLL(path)
        vp, i, d, l = C.c_void_p, C.c_int, C.c_double, C.c_long
        L.oref_num_threads.restype = i
        L.oref_set_num_threads.argtypes = [i]
        L.oref_generate_q.argtypes = [i, i, i, d, d, d, vp, l, l, vp]
        L.oref_matvec.argtypes = [i, i, i, d, d, d, vp, l, l, vp, vp, d, d, vp, C.POINTER(d)]
        L.oref_solver_cg.argtypes = [i, i, i, d, d, d, vp, l, l, vp, vp, d, l, d, vp, C.POINTER(l), C.POINTER(d)]
        L.oref_learn.argtypes = [i, i, i, d, d, d, d, vp, vp, l, l, vp, C.POINTER(d), C.POINTER(l), C.POINTER(d)]
        L.oref_predict.argtypes = [i, i, i, d, d, vp, vp, l, l, d, vp, l, vp]
        L.oref_kernel_function.argtypes = [i, i, i, d, d, vp, vp, l, C.POINTER(d)]
        _LIB = L
    return _LIB


def _dt(a: np.ndarray) -> int:
    if a.dtype == np.float32:
        return 0
    if a.dtype == np.float64:
        return 1
    raise TypeError(a.dtype)


def _p(a: np.ndarray):
    return a.ctypes.data_as(C.c_void_p)


def _chk(rc: int, what: str):
    if rc != 0:
        raise RuntimeError(f"oracle/_ref {what} failed (rc={rc})")


def num_threads() -> int:
    return lib().oref_num_threads()


def set_num_threads(t: int) -> None:
    lib().oref_set_num_threads(int(t))


def generate_q(X, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    q = np.empty(n - 1, X.dtype)
    _chk(lib().oref_generate_q(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, _p(X), n, F, _p(q)), "generate_q")
    return q


def matvec(X, q, d, QA_cost, add, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, ret=None, timing=None):
    """ret += add * Abar * d (run_device_kernel). Returns the updated ret."""
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    q = np.ascontiguousarray(q, X.dtype)
    d = np.ascontiguousarray(d, X.dtype)
    out = np.zeros(n - 1, X.dtype) if ret is None else np.array(ret, X.dtype, copy=True)
    secs = C.c_double(0)
    _chk(lib().oref_matvec(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, _p(X), n, F, _p(q), _p(d), QA_cost, add, _p(out), C.byref(secs)), "matvec")
    if timing is not None:
        timing["seconds"] = secs.value
    return out


def solver_cg(X, b, q, QA_cost, imax, eps, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    b = np.ascontiguousarray(b, X.dtype)
    q = np.ascontiguousarray(q, X.dtype)
    alpha = np.empty(n - 1, X.dtype)
    it = C.c_long(0)
    secs = C.c_double(0)
    _chk(lib().oref_solver_cg(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, _p(X), n, F, _p(b), _p(q), QA_cost, imax, eps, _p(alpha), C.byref(it), C.byref(secs)), "solver_cg")
    return alpha, it.value, secs.value


def learn(X, y, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, eps=1e-3):
    """Full csvm::learn(): returns (alpha[n], rho, iterations, seconds)."""
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    y = np.ascontiguousarray(y, X.dtype)
    alpha = np.empty(n, X.dtype)
    rho = C.c_double(0)
    it = C.c_long(0)
    secs = C.c_double(0)
    _chk(lib().oref_learn(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, eps, _p(X), _p(y), n, F, _p(alpha), C.byref(rho), C.byref(it), C.byref(secs)), "learn")
    return alpha, rho.value, it.value, secs.value


def predict(X, alpha, rho, points, kernel, degree=3, gamma=None, coef0=0.0):
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    alpha = np.ascontiguousarray(alpha, X.dtype)
    points = np.ascontiguousarray(points, X.dtype)
    out = np.empty(points.shape[0], X.dtype)
    _chk(lib().oref_predict(_dt(X), KERNELS[kernel], degree, gamma, coef0, _p(X), _p(alpha), n, F, rho, _p(points), points.shape[0], _p(out)), "predict")
    return out


def kernel_function(a, b, kernel, degree=3, gamma=1.0, coef0=0.0):
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b, a.dtype)
    out = C.c_double(0)
    _chk(lib().oref_kernel_function(_dt(a), KERNELS[kernel], degree, gamma, coef0, _p(a), _p(b), a.shape[0], C.byref(out)), "kernel_function")
    return out.value


# ---- file formats through the reference's own parsers / writer -------------------------------------------------------
def _io_sigs():
    L = lib()
    if getattr(L, "_io_ready", False):
        return L
    vp, i, d, l = C.c_void_p, C.c_int, C.c_double, C.c_long
    pl, pi, pd = C.POINTER(l), C.POINTER(i), C.POINTER(d)
    L.oref_parse_data_file.argtypes = [i, C.c_char_p, pl, pl, pi, vp, vp]
    L.oref_parse_model_file.argtypes = [i, C.c_char_p, pi, pi, pd, pd, pd, pl, pl, vp, vp, vp]
    L.oref_write_model.argtypes = [i, i, i, d, d, vp, vp, vp, l, l, d, C.c_char_p]
    L._io_ready = True
    return L


def parse_data_file(path: str, dtype=np.float64):
    """plssvm::parameter<T>::parse_file (src/plssvm/parameter.cpp:122-128). Returns (X, y or None)."""
    L = _io_sigs()
    dt = 0 if np.dtype(dtype) == np.float32 else 1
    n, F, has = C.c_long(0), C.c_long(0), C.c_int(0)
    _chk(L.oref_parse_data_file(dt, path.encode(), C.byref(n), C.byref(F), C.byref(has), None, None), "parse_data_file")
    X = np.empty((n.value, F.value), dtype)
    y = np.empty(n.value, dtype)
    _chk(L.oref_parse_data_file(dt, path.encode(), C.byref(n), C.byref(F), C.byref(has), _p(X), _p(y)), "parse_data_file")
    return X, (y if has.value else None)


def parse_model_file(path: str, dtype=np.float64):
    """plssvm::parameter<T>::parse_model_file (src/plssvm/parameter.cpp:366-520)."""
    L = _io_sigs()
    dt = 0 if np.dtype(dtype) == np.float32 else 1
    k, deg = C.c_int(0), C.c_int(0)
    g, c0, rho = C.c_double(0), C.c_double(0), C.c_double(0)
    n, F = C.c_long(0), C.c_long(0)
    args = (C.byref(k), C.byref(deg), C.byref(g), C.byref(c0), C.byref(rho), C.byref(n), C.byref(F))
    _chk(L.oref_parse_model_file(dt, path.encode(), *args, None, None, None), "parse_model_file")
    X = np.empty((n.value, F.value), dtype)
    alpha = np.empty(n.value, dtype)
    y = np.empty(n.value, dtype)
    _chk(L.oref_parse_model_file(dt, path.encode(), *args, _p(X), _p(alpha), _p(y)), "parse_model_file")
    kernel = {v: kk for kk, v in KERNELS.items()}[k.value]
    return dict(kernel=kernel, degree=deg.value, gamma=g.value, coef0=c0.value, rho=rho.value, X=X, alpha=alpha, y=y)


def write_model(path: str, X, y, alpha, rho, kernel, degree=3, gamma=None, coef0=0.0):
    """plssvm::csvm<T>::write_model (src/plssvm/csvm.cpp:60-204)."""
    L = _io_sigs()
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    y = np.ascontiguousarray(y, X.dtype)
    alpha = np.ascontiguousarray(alpha, X.dtype)
    _chk(L.oref_write_model(_dt(X), KERNELS[kernel], degree, gamma, coef0, _p(X), _p(y), _p(alpha), n, F, rho, path.encode()), "write_model")
