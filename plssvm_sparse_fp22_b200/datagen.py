"""Seeded synthetic inputs for parity tests and bench.py.

The reference's generator (utility_scripts/generate_data.py) is unseeded, dense-only and fails on this image's
scikit-learn (its ``feature_range=[-1, 1]`` list is rejected, utility_scripts/generate_data.py:75), so its math is
restated here with numpy only and an explicit seed:

* ``blobs``  — ``make_blobs(centers=2)`` (generate_data.py:54-56): two isotropic Gaussians, centres ~ U(-10, 10)^F,
  sigma = 1, samples split evenly, shuffled; labels ``2*y - 1`` (:73); per-feature min-max scaling to [-1, 1] (:75).
* ``planes`` — ``make_classification(n_informative=2, n_redundant=0, n_clusters_per_class=1)`` (:60-62): two informative
  features (class centroids at hypercube vertices, class_sep = 1, random linear mix), F-2 pure-noise features, 1 % label
  flips; then the same label map and scaling.
* sparse CSR — not in the reference: row nnz ~ Binomial(F, density), sorted unique columns,
  values U(-1, 1), labels = sign(x . w* + 0.1 N(0,1)).
"""
from __future__ import annotations

import numpy as np


def _minmax_scale(X: np.ndarray) -> np.ndarray:
    lo = X.min(axis=0)
    hi = X.max(axis=0)
    span = np.where(hi > lo, hi - lo, 1.0)
    X -= lo
    X *= 2.0 / span
    X -= 1.0
    return X


def make_dense(n: int, F: int, seed: int = 42, problem: str = "blobs", dtype=np.float64):
    """Returns (X[n, F], y[n]) with y in {-1, +1}; X scaled to [-1, 1] per feature."""
    rng = np.random.default_rng(seed)
    if problem == "blobs":
        centers = rng.uniform(-10.0, 10.0, size=(2, F))
        y01 = np.arange(n) % 2
        X = rng.standard_normal((n, F))
        X += centers[y01]
    elif problem == "planes":
        y01 = np.arange(n) % 2
        X = rng.standard_normal((n, F))
        centroids = np.array([[-1.0, -1.0], [1.0, 1.0]]) if F >= 2 else np.array([[-1.0], [1.0]])
        k = centroids.shape[1]
        A = 2.0 * rng.uniform(size=(k, k)) - 1.0
        X[:, :k] = X[:, :k] @ A + centroids[y01]
        flip = rng.uniform(size=n) < 0.01
        y01 = np.where(flip, 1 - y01, y01)
    else:
        raise ValueError(f"unknown problem {problem!r}")
    perm = rng.permutation(n)
    X = X[perm]
    y = (2 * y01[perm] - 1).astype(dtype)
    X = _minmax_scale(X).astype(dtype)
    return np.ascontiguousarray(X), y


def make_sparse_csr(n: int, F: int, density: float, seed: int = 42, dtype=np.float64, min_nnz: int = 1):
    """Returns (rowptr int64[n+1], colidx int32[nnz], values[nnz], y[n]).  Columns sorted and unique within a row."""
    rng = np.random.default_rng(seed)
    counts = rng.binomial(F, density, size=n).astype(np.int64)
    np.maximum(counts, min_nnz, out=counts)
    rowptr = np.zeros(n + 1, np.int64)
    np.cumsum(counts, out=rowptr[1:])
    nnz = int(rowptr[-1])
    # Draw columns row-block-wise: sample with replacement, then repair duplicates by re-drawing (density is small).
    colidx = np.empty(nnz, np.int32)
    rows = np.repeat(np.arange(n, dtype=np.int64), counts)
    cols = rng.integers(0, F, size=nnz, dtype=np.int64)
    for _ in range(64):
        key = rows * F + cols
        order = np.argsort(key, kind="stable")
        sk = key[order]
        dup = np.zeros(nnz, bool)
        dup[order[1:]] = sk[1:] == sk[:-1]
        ndup = int(dup.sum())
        if ndup == 0:
            break
        cols[dup] = rng.integers(0, F, size=ndup, dtype=np.int64)
    else:
        raise RuntimeError("could not de-duplicate columns (density too high for this generator)")
    key = rows * F + cols
    order = np.argsort(key, kind="stable")
    colidx[:] = cols[order].astype(np.int32)
    values = rng.uniform(-1.0, 1.0, size=nnz)
    wstar = rng.standard_normal(F)
    score = np.add.reduceat(values * wstar[colidx], rowptr[:-1]) if nnz else np.zeros(n)
    score = score + 0.1 * rng.standard_normal(n)
    y = np.where(score > 0, 1.0, -1.0).astype(dtype)
    return rowptr, colidx, values.astype(dtype), y


def csr_to_dense(rowptr, colidx, values, F: int) -> np.ndarray:
    n = rowptr.shape[0] - 1
    X = np.zeros((n, F), values.dtype)
    rows = np.repeat(np.arange(n), np.diff(rowptr))
    X[rows, colidx] = values
    return X


def dense_to_csr(X: np.ndarray):
    n, F = X.shape
    mask = X != 0
    counts = mask.sum(axis=1).astype(np.int64)
    rowptr = np.zeros(n + 1, np.int64)
    np.cumsum(counts, out=rowptr[1:])
    rows, cols = np.nonzero(mask)
    return rowptr, cols.astype(np.int32), np.ascontiguousarray(X[rows, cols])


def write_libsvm(path: str, X: np.ndarray, y=None, one_based: bool = False) -> None:
    """LIBSVM text, indices as sklearn's dump_svmlight_file writes them by default (0-based; generate_data.py:94-95),
    zeros skipped."""
    off = 1 if one_based else 0
    with open(path, "w") as f:
        for i in range(X.shape[0]):
            nz = np.nonzero(X[i])[0]
            feats = " ".join(f"{j + off}:{float(X[i, j])!r}" if X.dtype == np.float64 else f"{j + off}:{float(X[i, j]):.9g}" for j in nz)
            if y is not None:
                lab = int(y[i]) if float(y[i]).is_integer() else float(y[i])
                f.write(f"{lab} {feats}\n")
            else:
                f.write(f"{feats}\n")


# --- FP22 (1 sign | 8 exponent | 13 mantissa = top 22 bits of binary32, round-to-nearest-even) host codec -----------
def fp22_encode(x: np.ndarray) -> np.ndarray:
    u = np.ascontiguousarray(x, np.float32).view(np.uint32).astype(np.uint64)
    special = (u & 0x7F800000) == 0x7F800000
    lsb = (u >> 10) & 1
    r = ((u + 0x1FF + lsb) >> 10).astype(np.uint32)
    t = (u >> 10).astype(np.uint32)
    nan_fix = special & ((u & 0x007FFFFF) != 0) & ((t & 0x1FFF) == 0)
    t = np.where(nan_fix, t | 1, t)
    return np.where(special, t, r).astype(np.uint32)


def fp22_decode(v22: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(v22, np.uint32) << np.uint32(10)).view(np.float32)


def fp22_round(x: np.ndarray) -> np.ndarray:
    return fp22_decode(fp22_encode(x))


def pack_fp22_csr(rowptr: np.ndarray, colidx: np.ndarray, values: np.ndarray):
    """CSR -> FP22-packed CSR: one uint32 per stored entry, ``(fp22(value) << 10) | column_delta`` with col_0 = delta_0 and
    col_k = col_{k-1} + delta_k; gaps > 1023 are bridged by padding entries (value +0, delta 1023).
    Returns (rowptr_packed int64[n+1], packed uint32[...])."""
    rowptr = np.ascontiguousarray(rowptr, np.int64)
    n = rowptr.shape[0] - 1
    nnz = int(rowptr[-1])
    if nnz == 0:
        return rowptr.copy(), np.zeros(0, np.uint32)
    cols = np.ascontiguousarray(colidx, np.int64)
    prev = np.empty(nnz, np.int64)
    prev[1:] = cols[:-1]
    starts = rowptr[:-1][np.diff(rowptr) > 0]
    prev[starts] = 0
    delta = cols - prev
    npad = np.maximum((delta - 1) // 1023, 0)
    last = delta - 1023 * npad
    code = fp22_encode(np.asarray(values, np.float32)).astype(np.uint32)
    out_counts = npad + 1
    total = int(out_counts.sum())
    packed = np.full(total, np.uint32(1023), np.uint32)  # padding: value +0, delta 1023
    pos = np.cumsum(out_counts) - 1  # position of the real entry of each nnz
    packed[pos] = (code << np.uint32(10)) | last.astype(np.uint32)
    rows = np.repeat(np.arange(n, dtype=np.int64), np.diff(rowptr))
    per_row = np.bincount(rows, weights=out_counts, minlength=n).astype(np.int64)
    rp = np.zeros(n + 1, np.int64)
    np.cumsum(per_row, out=rp[1:])
    return rp, packed


def unpack_fp22_csr(rowptr_packed: np.ndarray, packed: np.ndarray):
    """Inverse of pack_fp22_csr (padding entries dropped): (rowptr, colidx int32, values float32)."""
    n = rowptr_packed.shape[0] - 1
    rp = np.zeros(n + 1, np.int64)
    cols, vals = [], []
    for i in range(n):
        w = packed[rowptr_packed[i]:rowptr_packed[i + 1]]
        c = np.cumsum((w & np.uint32(1023)).astype(np.int64))
        v = (w & np.uint32(0xFFFFFC00)).view(np.float32)
        keep = (w & np.uint32(0x7FFFFC00)) != 0
        cols.append(c[keep].astype(np.int32))
        vals.append(v[keep])
        rp[i + 1] = rp[i] + int(keep.sum())
    return rp, (np.concatenate(cols) if cols else np.zeros(0, np.int32)), (np.concatenate(vals) if vals else np.zeros(0, np.float32))
