"""BASELINE INFRASTRUCTURE — ctypes binding of oracle/_ref/liboracle_ref_cuda.so: the UNMODIFIED reference CUDA backend
(src/plssvm/backends/CUDA/*, backends/gpu_csvm.cpp) compiled for sm_100 by ``make -C oracle ref_cuda``.  SURVEY.md §8 row a10
calls these "the kernels to beat": bench.py times them on the same B200 next to the repo's kernels.  NOT product code."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None
KERNELS = {"linear": 0, "polynomial": 1, "rbf": 2}


def lib_path() -> str:
    return os.path.join(_HERE, "_ref", "liboracle_ref_cuda.so")


def available() -> bool:
    return os.path.exists(lib_path())


def lib():
    global _LIB
    if _LIB is None:
        L = C.CDLL(lib_path())
        vp, i, d, l = C.c_void_p, C.c_int, C.c_double, C.c_long
        L.orefcuda_solver_cg.argtypes = [i, i, i, d, d, d, vp, l, l, vp, vp, d, l, d, vp, C.POINTER(l), C.POINTER(d), C.POINTER(d)]
        L.orefcuda_matvec_ms.argtypes = [i, i, i, d, d, d, vp, l, l, vp, vp, d, d, vp, i, C.POINTER(d)]
        _LIB = L
    return _LIB


def _dt(a):
    return 0 if a.dtype == np.float32 else 1


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def matvec_ms(X, q, d, QA_cost, add, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0, reps=3):
    """ret = add * Abar d by the reference's device_kernel_{linear,poly,radial}; returns (ret, mean ms, best ms) of `reps` more launches."""
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    q = np.ascontiguousarray(q, X.dtype)
    d = np.ascontiguousarray(d, X.dtype)
    out = np.zeros(n - 1, X.dtype)
    ms = (C.c_double * 2)()
    rc = lib().orefcuda_matvec_ms(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, _p(X), n, F, _p(q), _p(d), QA_cost, add, _p(out), reps, ms)
    if rc != 0:
        raise RuntimeError(f"reference CUDA backend: matvec failed (rc={rc})")
    return out, ms[0], ms[1]


def solver_cg(X, b, q, QA_cost, imax, eps, kernel, degree=3, gamma=None, coef0=0.0, cost=1.0):
    """gpu_csvm::solver_CG of the reference; returns (alpha, iterations, seconds of solver_CG, seconds of setup_data_on_device)."""
    X = np.ascontiguousarray(X)
    n, F = X.shape
    gamma = 1.0 / F if gamma is None else gamma
    b = np.ascontiguousarray(b, X.dtype)
    q = np.ascontiguousarray(q, X.dtype)
    alpha = np.empty(n - 1, X.dtype)
    it, secs, setup = C.c_long(0), C.c_double(0), C.c_double(0)
    rc = lib().orefcuda_solver_cg(_dt(X), KERNELS[kernel], degree, gamma, coef0, cost, _p(X), n, F, _p(b), _p(q), QA_cost, imax, eps, _p(alpha),
                                  C.byref(it), C.byref(secs), C.byref(setup))
    if rc != 0:
        raise RuntimeError(f"reference CUDA backend: solver_CG failed (rc={rc})")
    return alpha, it.value, secs.value, setup.value
