"""Sparse-path timing on one B200 (design input): ms per implicit matvec, Gram products/s, kernel evaluations/s.
usage: python tools/bench_sparse.py --points N --features F --density RHO --kernel rbf|polynomial --dtype f64|f32|fp22 [--block W] [--reps K]
"""
import argparse
import json
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from plssvm_sparse_fp22_b200 import capi, datagen  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--points", type=int, default=131072)
    ap.add_argument("--features", type=int, default=100000)
    ap.add_argument("--density", type=float, default=1e-3)
    ap.add_argument("--kernel", default="rbf")
    ap.add_argument("--dtype", default="f64", choices=["f64", "f32", "fp22"])
    ap.add_argument("--block", type=int, default=0)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--lanes", type=int, default=0)
    ap.add_argument("--threads", type=int, default=0)
    ap.add_argument("--group", type=int, default=0)
    ap.add_argument("--cache", type=int, default=0, help="0 auto, 1 off (on the fly), 2 required")
    a = ap.parse_args()
    dt = np.float64 if a.dtype == "f64" else np.float32
    t0 = time.time()
    rp, ci, v, y = datagen.make_sparse_csr(a.points, a.features, a.density, seed=45, dtype=dt)
    t_gen = time.time() - t0
    cf = np.bincount(ci[: rp[-2]], minlength=a.features).astype(np.float64)  # rows 0..m-1
    products = float((cf * cf).sum())
    with capi.Handle(dt, a.kernel, 3, 1.0 / a.features, 0.0, 1.0) as h:
        if a.block:
            h.set_option(capi.OPT_SPARSE_BLOCK, a.block)
        if a.lanes:
            h.set_option(capi.OPT_SPARSE_LANES, a.lanes)
        if a.threads:
            h.set_option(capi.OPT_SPARSE_THREADS, a.threads)
        if a.group:
            h.set_option(capi.OPT_SPARSE_GROUP, a.group)
        h.set_option(capi.OPT_SPARSE_CACHE, a.cache)
        t0 = time.time()
        if a.dtype == "fp22":
            prp, packed = datagen.pack_fp22_csr(rp, ci, v)
            h.set_data_csr(prp, None, packed, a.features, value_format=capi.VAL_FP22)
        else:
            h.set_data_csr(rp, ci, v, a.features)
        t_set = time.time() - t0
        q = h.generate_q()
        h.set_q(q, h.kxx_last() + 1.0)
        d = np.random.default_rng(1).standard_normal(a.points - 1).astype(dt)
        t0 = time.time()
        h.matvec(d)  # warm-up (assembles the S cache when enabled)
        t_first = time.time() - t0
        cache_entries = h.stats().sparse_cache_entries
        build_ms = h.stats().cache_build_ms
        h.reset_stats()
        for _ in range(a.reps):
            out = h.matvec(d)
        st = h.stats()
        ms = st.matvec_ms / max(st.matvecs_timed, 1)
    print(json.dumps({"points": a.points, "features": a.features, "density": a.density, "kernel": a.kernel, "dtype": a.dtype, "block": a.block, "lanes": a.lanes, "threads": a.threads, "group": a.group,
                      "nnz": int(rp[-1]), "gram_products": products, "ms_per_matvec": ms, "gproducts_per_s": products / ms / 1e6,
                      "pairs_per_s": (a.points - 1.0) ** 2 / ms / 1e6 * 1e-3, "cache": a.cache, "cache_entries": int(cache_entries), "first_matvec_s": t_first, "cache_build_ms": build_ms, "gen_s": t_gen, "set_data_s": t_set, "checksum": float(np.abs(out).sum())}))


if __name__ == "__main__":
    main()
