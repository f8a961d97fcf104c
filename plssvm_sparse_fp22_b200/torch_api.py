"""fit / predict on torch CUDA tensors — a thin convenience layer over the C ABI.

The dense entry points of libb200svm accept device pointers (include/b200svm.h), so tensors that already live in HBM are handed over
by ``data_ptr()``: no host round trip, the only copy is the library's own device-to-device staging into its padded layout.  All work
runs on the tensor's device and on torch's current stream.  There is no CPU fallback: CPU tensors are rejected.

    svm = LSSVM(kernel="rbf", cost=1.0).fit(X, y)          # X: (n, F) float32 / float64 CUDA tensor, y: (n,) labels +-1
    labels = svm.predict(points)                           # (P,) tensor of +-1 on the same device
    values = svm.decision_function(points)

Semantics are those of plssvm::csvm::learn / predict / predict_label (src/plssvm/csvm.cpp:207-267, 343-366): the last row is the
reference's "last point", alpha_n = -sum(alpha), label = value > 0 ? +1 : -1.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import capi


class LSSVM:
    def __init__(self, kernel="linear", degree=3, gamma=None, coef0=0.0, cost=1.0, epsilon=1e-3):
        self.kernel, self.degree, self.gamma, self.coef0, self.cost, self.epsilon = kernel, degree, gamma, coef0, cost, epsilon
        self.h = None
        self.alpha = None
        self.rho = None
        self.iterations = None

    @staticmethod
    def _check(t, name):
        import torch
        if not isinstance(t, torch.Tensor) or not t.is_cuda:
            raise TypeError(f"{name} must be a torch CUDA tensor (libb200svm has no CPU path)")
        if t.dtype not in (torch.float32, torch.float64):
            raise TypeError(f"{name} must be float32 or float64")
        return t.contiguous()

    def fit(self, X, y):
        import torch
        X = self._check(X, "X")
        y = self._check(y.to(X.dtype), "y")
        n, F = X.shape
        if y.shape != (n,):
            raise ValueError("y must have one label per row of X")
        np_dtype = np.float32 if X.dtype == torch.float32 else np.float64
        if self.h is not None:
            self.h.close()
        gamma = self.gamma if self.gamma is not None else 1.0 / F  # src/plssvm/parameter.cpp:150-152
        self.h = capi.Handle(np_dtype, self.kernel, self.degree, gamma, self.coef0, self.cost, device=X.device.index or 0)
        h, L = self.h, self.h.L
        with torch.cuda.device(X.device):
            h.set_stream(torch.cuda.current_stream().cuda_stream)
            h._chk(L.b200svm_set_data_dense(h.h, C.c_void_p(X.data_ptr()), n, F, X.stride(0)))
            h.n, h.F = n, F
            self.alpha = torch.empty(n, dtype=X.dtype, device=X.device)
            rho, it = C.c_double(0), C.c_int64(0)
            h._chk(L.b200svm_learn(h.h, C.c_void_p(y.data_ptr()), float(self.epsilon), C.c_void_p(self.alpha.data_ptr()), C.byref(rho), C.byref(it)))
        self.rho, self.iterations = rho.value, it.value
        return self

    def decision_function(self, points):
        import torch
        if self.h is None:
            raise RuntimeError("fit() first")
        points = self._check(points.to(self.alpha.dtype), "points")
        P, F = points.shape
        if F != self.h.F:
            raise ValueError("Number of features per data point must match the number of features per predict point!")
        out = torch.empty(P, dtype=points.dtype, device=points.device)
        with torch.cuda.device(points.device):
            self.h._chk(self.h.L.b200svm_predict_dense(self.h.h, C.c_void_p(self.alpha.data_ptr()), float(self.rho), C.c_void_p(points.data_ptr()), P,
                                                       points.stride(0), C.c_void_p(out.data_ptr())))
        return out

    def predict(self, points):
        import torch
        v = self.decision_function(points)
        return torch.where(v > 0, torch.ones_like(v), -torch.ones_like(v))  # operators::sign: 0 -> -1

    def close(self):
        if self.h is not None:
            self.h.close()
            self.h = None
