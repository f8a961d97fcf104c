"""Predict throughput (BASELINE.json configs[4]: sparse test points against a sparse RBF model), one process per GPU.

    python tools/bench_predict.py                                   # scaled-down single-GPU run
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/bench_predict.py \
           --svs 262144 --features 20000 --density 1e-2 --points 4194304        # configs[4]

The test rows are split evenly over the ranks (no collective on the data path: every rank holds the whole model); a rank draws its
own share of the points (seed + rank).  The timed region is one b200svm_predict_csr call per rank with HOST buffers — upload of the
points, decision values for all of them, download — i.e. what plssvm-predict does after parsing (src/plssvm/csvm.cpp:343-366 ->
backends/OpenMP/csvm.cpp:191-244).  value = all points / max over ranks of that time.  Prints one JSON line on rank 0.
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--svs", type=int, default=32768)
    ap.add_argument("--features", type=int, default=20000)
    ap.add_argument("--density", type=float, default=1e-2)
    ap.add_argument("--points", type=int, default=131072, help="test points over all ranks")
    ap.add_argument("--kernel", default="rbf")
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--storage", default="native", choices=["native", "fp22"])
    ap.add_argument("--reps", type=int, default=2)
    ap.add_argument("--seed", type=int, default=46)
    ap.add_argument("--sweep", default="", help="geometry sweep on one GPU: semicolon list of block,threads,lanes (0 = default), e.g. '0,0,0;16384,256,0'")
    a = ap.parse_args()

    import torch
    import torch.distributed as dist

    from plssvm_sparse_fp22_b200 import capi, datagen

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dt = np.float32 if a.dtype == "f32" else np.float64
    F = a.features
    t0 = time.time()
    rp, ci, v, _ = datagen.make_sparse_csr(a.svs, F, a.density, seed=a.seed, dtype=dt)
    share = a.points // world
    prp, pci, pv, _ = datagen.make_sparse_csr(share, F, a.density, seed=a.seed + 1 + rank, dtype=dt, min_nnz=0)
    fmt = capi.VAL_F32 if dt == np.float32 else capi.VAL_F64
    if a.storage == "fp22":
        assert dt == np.float32
        rp, v = datagen.pack_fp22_csr(rp, ci, v)
        prp, pv = datagen.pack_fp22_csr(prp, pci, pv)
        ci = pci = None
        fmt = capi.VAL_FP22
    t_gen = time.time() - t0
    rng = np.random.default_rng(a.seed)
    alpha = (rng.standard_normal(a.svs) * 0.1).astype(dt)  # alpha ~ N(0, 1e-2), rho = 0.1 (SURVEY.md §8d cfg 5)
    for cfg in [c for c in a.sweep.split(";") if c]:
        blk, thr, lanes = (int(x) for x in cfg.split(","))
        with capi.Handle(dt, a.kernel, 3, 1.0 / F, 0.0, 1.0, device=local) as h:
            if blk:
                h.set_option(capi.OPT_SPARSE_BLOCK, blk)
            if thr:
                h.set_option(capi.OPT_SPARSE_THREADS, thr)
            if lanes:
                h.set_option(capi.OPT_SPARSE_LANES, lanes)
            h.set_data_csr(rp, ci, v, F, value_format=fmt)
            h.predict_csr(alpha, 0.1, prp[:2], None if pci is None else pci[: prp[1]], pv[: prp[1]], value_format=fmt)
            ts = []
            for _ in range(a.reps):
                t0 = time.perf_counter()
                dec = h.predict_csr(alpha, 0.1, prp, pci, pv, value_format=fmt)
                ts.append(time.perf_counter() - t0)
            print(json.dumps({"sweep": cfg, "seconds": min(ts), "points_per_s": share / min(ts), "checksum": float(np.abs(dec).sum())}), flush=True)
    with capi.Handle(dt, a.kernel, 3, 1.0 / F, 0.0, 1.0, device=local) as h:
        h.set_data_csr(rp, ci, v, F, value_format=fmt)
        dec = h.predict_csr(alpha, 0.1, prp[:2], None if pci is None else pci[: prp[1]], pv[: prp[1]], value_format=fmt)  # warm-up (one point)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        times = []
        for _ in range(a.reps):
            t0 = time.perf_counter()
            dec = h.predict_csr(alpha, 0.1, prp, pci, pv, value_format=fmt)
            times.append(time.perf_counter() - t0)
        st = h.stats()
    t = min(times)
    labels_pos = int((dec > 0).sum())
    if world > 1:
        tt = torch.tensor([t], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t = float(tt[0])
        cnt = torch.tensor([labels_pos], device="cuda", dtype=torch.int64)
        dist.all_reduce(cnt)
        labels_pos = int(cnt[0])
    if rank == 0:
        total = share * world
        evals = float(total) * a.svs
        print(json.dumps({
            "metric": "predict_points_per_sec", "value": total / t, "unit": "points/s", "n_gpus": world, "seconds": t,
            "kernel_evals_per_s": evals / t, "gram_products_per_s_est": evals * F * a.density * a.density / t,
            "config": {"workload": f"predict: {total} sparse test points x {a.svs} support vectors, {F} features at density {a.density:g}, "
                                   f"{a.kernel}, {a.dtype}{' FP22 rows' if a.storage == 'fp22' else ''}",
                       "points": total, "svs": a.svs, "features": F, "density": a.density},
            "timed": "b200svm_predict_csr with host buffers (H2D of the points, kernels, D2H of the decision values), wall clock, best of %d" % a.reps,
            "h2d_bytes": int(st.h2d_bytes), "gpu_launches": int(st.kernel_launches), "positive_labels": labels_pos, "datagen_s": t_gen}), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
