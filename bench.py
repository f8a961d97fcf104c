#!/usr/bin/env python
"""bench.py — CG iterations/sec of the LS-SVM hot path (BASELINE.json metric) on N B200s of one node.

A "step" is one CG iteration (implicit matvec Abar*d + the fused vector updates / dot products) on the BASELINE
configs[1] workload: RBF kernel, 65536 points x 1024 dense features, float, gamma = 1/F, C = 1 — run in fixed-iteration
mode (eps = 0) because at the configured eps = 1e-3 CG converges after 1-2 iterations.  A timed solve of K
steps therefore contains K + 1 applications of the implicit matrix (the initial residual, csvm.cpp:94) plus one more per
50 iterations (csvm.cpp:130-135) — they are inside the timed region, nothing is skipped.

  value : K / t, t = device time (CUDA events on the launching stream, max over ranks) of one b200svm_cg_solve call with
          the training matrix already resident in HBM.
  e2e   : K / t_e2e through the C ABI with HOST buffers: b200svm_set_data_dense (H2D of X from pinned memory) +
          b200svm_generate_q + b200svm_cg_solve (H2D of b, D2H of alpha) inside the timed region.
  roofline : the O(m^2 F) Gram-tile kernel; algorithmic flops per launch = F * m * (m+1), duration =
          CUDA events recorded around each launch inside libb200svm (b200svm_stats.matvec_ms / matvecs).
  cpu_baseline / --impl reference : the UNMODIFIED reference OpenMP backend (oracle/_ref, built from /root/reference) or,
          if absent, the oracle port, timed on the host cores on a bounded row-subset of the same workload and scaled by
          the pair count to one full CG iteration.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# NCCL prints "NCCL version ..." on stdout at NCCL_DEBUG=VERSION: keep stdout to the one JSON line (INFO / TRACE are left alone)
if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
    os.environ["NCCL_DEBUG"] = "WARN"

METRIC = "cg_iterations_per_sec"
UNIT = "iterations/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--points", type=int, default=65536)
    ap.add_argument("--features", type=int, default=1024)
    ap.add_argument("--kernel", default="rbf", choices=["linear", "polynomial", "rbf"])
    ap.add_argument("--dtype", default="f32", choices=["f32", "f64"])
    ap.add_argument("--path", type=int, default=0, help="matvec path: 0 auto, 1 SIMT, 2 tcgen05")
    ap.add_argument("--cpu-sample-points", type=int, default=0, help="row subset for the CPU baseline (0 = calibrate to ~10 s)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--seed", type=int, default=43)
    ap.add_argument("--density", type=float, default=0.0,
                    help="> 0: SPARSE workload (not the driver's default line): CSR rows of this density, e.g. configs[2] "
                         "--points 1000000 --features 100000 --density 1e-3 --dtype f64, configs[3] --points 262144 --features 20000 "
                         "--density 1e-2 --kernel polynomial --storage fp22")
    ap.add_argument("--storage", default="native", choices=["native", "fp22"], help="sparse value storage")
    ap.add_argument("--dense-cache", type=int, default=0,
                    help="dense tcgen05 path: 0 store the kernel-matrix tiles on first use if they fit (matvec = HBM-bound sweep), "
                         "1 recompute the Gram tiles on the tensor cores in every matvec")
    ap.add_argument("--peer-exchange", type=int, default=0, help="multi-GPU: 0 peer-memory exchange (default), 1 ncclAllReduce")
    ap.add_argument("--skip-e2e", action="store_true", help="large sparse runs only: skip the host-buffer leg (e2e = null)")
    ap.add_argument("--cache", type=int, default=0, help="sparse: 0 cache S if it fits, 1 recompute on the fly, 2 require the cache")
    return ap.parse_args()


def workload_config(a):
    if a.density > 0:
        return {
            "workload": f"sparse: {a.kernel} kernel, {a.points} points x {a.features} features at density {a.density:g} (CSR, "
                        f"{'FP22-packed' if a.storage == 'fp22' else a.dtype} values), gamma=1/F, C=1, fixed-iteration CG (eps=0)",
            "points": a.points, "features": a.features, "kernel": a.kernel, "density": a.density, "storage": a.storage,
            "l2": "inputs larger than L2 (the cached S matrix / the transpose are streamed from HBM every step)",
        }
    return {
        "workload": f"configs[1]: {a.kernel} kernel, {a.points} points x {a.features} dense features, {a.dtype}, gamma=1/F, C=1, "
                    "fixed-iteration CG (eps=0)",
        "points": a.points, "features": a.features, "kernel": a.kernel,
        "l2": "inputs larger than L2: every step streams the stored kernel-matrix tiles (8.6 GB at 65536 points) or, with --dense-cache 1, "
              "re-reads the operand planes of X (%.0f MB) for every tile; no flush needed" % (2 * a.points * a.features * (4 if a.dtype == "f32" else 8) / 1e6),
        "dense_cache": "auto (kernel-matrix tiles stored once inside the timed region)" if a.dense_cache == 0 else "off (Gram tiles recomputed every matvec)",
    }


def alg_flops_per_matvec(n, F):
    m = n - 1
    return float(F) * m * (m + 1)  # 2 F P, P = m(m+1)/2 unordered pairs incl. diagonal


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: reference OpenMP backend (oracle/_ref) or the oracle port, on a bounded sample
# ----------------------------------------------------------------------------------------------------------------------
def cpu_reference_iteration_rate(a, steps, warmup, X, sample_points=0, budget_s=10.0):
    """Times the reference's run_device_kernel on the first n_s points (all F features) and scales by the pair count."""
    from oracle import port, ref
    dt = np.float32 if a.dtype == "f32" else np.float64
    use_ref = ref.available()
    cores = os.cpu_count() or 1
    if use_ref:
        ref.set_num_threads(cores)

    def one(ns):
        Xs = np.ascontiguousarray(X[:ns].astype(dt))
        d = np.ones(ns - 1, dt)
        q = np.zeros(ns - 1, dt)
        t0 = time.perf_counter()
        if use_ref:
            ref.matvec(Xs, q, d, 1.0, 1.0, a.kernel)
        else:
            port.matvec(Xs, q, d, 1.0, 1.0, a.kernel, rows=(0, ns - 1))
        return time.perf_counter() - t0

    ns = sample_points
    if ns <= 0:
        t_probe = one(1024)
        per_pair = t_probe / (1023 * 1024 / 2 if use_ref else 1023 * 1023)
        ns = int(min(a.points, max(1024, (2 * (budget_s / max(steps + warmup, 1)) / per_pair) ** 0.5 if use_ref
                                   else (budget_s / max(steps + warmup, 1) / per_pair) ** 0.5)))
        ns = max(1024, ns // 256 * 256)
    for _ in range(warmup):
        one(ns)
    ts = [one(ns) for _ in range(steps)]
    t = float(np.sum(ts))
    m_s, m = ns - 1, a.points - 1
    pairs_s = m_s * (m_s + 1) / 2 if use_ref else float(m_s) * m_s
    pairs_full = m * (m + 1) / 2 if use_ref else float(m) * m
    t_iter_full = (t / steps) * pairs_full / pairs_s  # one CG iteration = one matvec (+ O(m) vector ops, ignored)
    return {
        "value": 1.0 / t_iter_full, "unit": UNIT, "cores": cores, "kind": "reference" if use_ref else "port",
        "sample": f"run_device_kernel on the first {ns} of {a.points} points (all {a.features} features), {steps} timed calls "
                  f"({t:.1f} s CPU wall), scaled by the pair count to one full iteration",
        "ms_per_step": 1e3 * t_iter_full, "seconds_measured": t,
    }


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from plssvm_sparse_fp22_b200 import datagen
    X, _ = datagen.make_dense(min(a.points, 16384), a.features, seed=a.seed, problem="blobs", dtype=np.float32 if a.dtype == "f32" else np.float64)
    res = cpu_reference_iteration_rate(a, a.steps, a.warmup, X, a.cpu_sample_points, budget_s=60.0)
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic", "config": workload_config(a), "impl": "reference",
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed regions.  NVML in-process (a sample every ~5 ms: the default timed region is
    only tens of milliseconds long); `nvidia-smi` (one sample per ~100 ms) if NVML cannot be initialised."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []   # (sm_mhz, sm_max_mhz, [active reason names])
        self.stop_flag = threading.Event()
        self.source = "nvidia-smi"
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = index
            if vis:
                tok = vis.split(",")[index].strip()
                phys = int(tok) if tok.isdigit() else None
            self.dev = pynvml.nvmlDeviceGetHandleByIndex(phys) if phys is not None else pynvml.nvmlDeviceGetHandleByUUID(tok.encode())
            self.masks = [(nm, getattr(pynvml, c)) for nm, c in zip(self.NAMES, ("nvmlClocksEventReasonHwSlowdown", "nvmlClocksEventReasonHwThermalSlowdown",
                                                                                "nvmlClocksEventReasonSwThermalSlowdown", "nvmlClocksEventReasonSwPowerCap"))
                          if hasattr(pynvml, c)]
            if len(self.masks) != 4:
                self.masks = [(nm, getattr(pynvml, c)) for nm, c in zip(self.NAMES, ("nvmlClocksThrottleReasonHwSlowdown", "nvmlClocksThrottleReasonHwThermalSlowdown",
                                                                                    "nvmlClocksThrottleReasonSwThermalSlowdown", "nvmlClocksThrottleReasonSwPowerCap"))]
            self.sm_max = float(pynvml.nvmlDeviceGetMaxClockInfo(self.dev, pynvml.NVML_CLOCK_SM))
            pynvml.nvmlDeviceGetClockInfo(self.dev, pynvml.NVML_CLOCK_SM)  # probe
            self.nv = pynvml
            self.source = "nvml"
        except Exception:
            self.nv = None

    def run(self):
        while not self.stop_flag.is_set():
            try:
                if self.nv is not None:
                    sm = float(self.nv.nvmlDeviceGetClockInfo(self.dev, self.nv.NVML_CLOCK_SM))
                    try:
                        bits = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev))
                    except Exception:
                        bits = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev))
                    self.samples.append((sm, self.sm_max, [nm for nm, m in self.masks if bits & m]))
                    self.stop_flag.wait(0.005)
                    continue
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-i", str(self.index)],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    f = [t.strip() for t in out.split(",")]
                    if len(f) >= 7 and f[0].replace(".", "").isdigit() and f[1].replace(".", "").isdigit():
                        self.samples.append((float(f[0]), float(f[1]), [nm for k, nm in enumerate(self.NAMES) if f[3 + k].lower().startswith("active")]))
            except Exception:
                pass
            self.stop_flag.wait(0.1)

    def summary(self):
        sm = [s[0] for s in self.samples]
        mx = [s[1] for s in self.samples]
        reasons = [nm for nm in self.NAMES if any(nm in s[2] for s in self.samples)]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(self.samples), "source": self.source}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d, "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback"


def run_b200_arm(a):
    import torch
    import torch.distributed as dist

    from plssvm_sparse_fp22_b200 import capi, datagen

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dt = np.float32 if a.dtype == "f32" else np.float64
    n, F, K, W = a.points, a.features, a.steps, a.warmup

    sparse = a.density > 0
    if sparse:
        rp, ci, vals, y = datagen.make_sparse_csr(n, F, a.density, seed=a.seed, dtype=dt)
        if a.storage == "fp22":
            assert a.dtype == "f32", "FP22 storage needs --dtype f32"
            prp, packed = datagen.pack_fp22_csr(rp, ci, vals)
        X = None
    else:
        X, y = datagen.make_dense(n, F, seed=a.seed, problem="blobs", dtype=dt)
        # pinned host staging for the e2e leg
        Xp = torch.empty((n, F), dtype=torch.float32 if dt == np.float32 else torch.float64, pin_memory=True)
        Xp.numpy()[:] = X
        Xh = Xp.numpy()
    b = (y[:-1] - y[-1]).astype(dt)

    def set_data():
        if not sparse:
            h.set_data(Xh)
        elif a.storage == "fp22":
            h.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
        else:
            h.set_data_csr(rp, ci, vals, F)

    h = capi.Handle(dt, a.kernel, 3, 1.0 / F, 0.0, 1.0, device=local_rank)
    stream = torch.cuda.current_stream()
    h.set_stream(stream.cuda_stream)
    if a.path:
        h.set_option(capi.OPT_MATVEC_PATH, a.path)
    if sparse:
        h.set_option(capi.OPT_SPARSE_CACHE, a.cache)
    else:
        h.set_option(capi.OPT_DENSE_CACHE, a.dense_cache)

    def drop_caches():
        # derived data (cached kernel-matrix tiles / sparse S) is NOT carried into a timed region: its assembly is part of the job
        if sparse:
            h.set_option(capi.OPT_SPARSE_CACHE, a.cache)
        else:
            h.set_option(capi.OPT_DENSE_CACHE, a.dense_cache)
    if world > 1:
        h.set_option(capi.OPT_PEER_EXCHANGE, a.peer_exchange)
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idbuf.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idbuf, 0)
        h.comm_init(bytes(idbuf.cpu().numpy().tobytes()), rank, world)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- resident leg (value) ----
    set_data()
    q = h.generate_q()
    QA = h.kxx_last() + 1.0
    # warm-up (untimed): solves of max(W, 3) steps, repeated until the device has been busy for ~0.3 s — the first solves of a fresh
    # process run the same kernels up to 1.4x slower (tools/probe_sweep.py) and a 3-step solve lasts only ~13 ms.  The repeat count
    # is taken from the first solve and agreed over the ranks (every solve contains exchanges).
    t_warm = time.perf_counter()
    h.solver_cg(b, None, QA, max(W, 3), 0.0)
    reps = int(min(50, max(1, 0.3 / max(time.perf_counter() - t_warm, 1e-4))))
    if world > 1:
        rt = torch.tensor([reps], device="cuda", dtype=torch.int64)
        dist.all_reduce(rt, op=dist.ReduceOp.MAX)
        reps = int(rt[0])
    for _ in range(reps):
        h.solver_cg(b, None, QA, max(W, 3), 0.0)
    warm_steps = (1 + reps) * max(W, 3)
    drop_caches()
    barrier()
    sampler = ClockSampler(local_rank)
    sampler.start()
    h.reset_stats()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    alpha, iters, delta = h.solver_cg(b, None, QA, K, 0.0)
    e1.record(stream)
    barrier()
    t_ms = e0.elapsed_time(e1)
    st = h.stats()
    assert iters == K, (iters, K)
    launches = int(st.kernel_launches)
    build_ms = float(st.cache_build_ms)
    dense_cache_bytes = int(st.dense_cache_bytes)
    # the matvec that assembles a cache is timed as a whole: take the assembly out to get the steady-state launch duration
    n_matvecs = int(st.matvecs_timed)
    if not sparse and dense_cache_bytes > 0 and build_ms > 0:
        n_matvecs -= 1  # dense tiles: the STORE pass of the tcgen05 kernel also serves the matvec that triggered it (no sweep in that one)
    matvec_ms = (st.matvec_ms - build_ms) / max(n_matvecs, 1)
    path = int(st.matvec_path)
    cache_entries = int(st.sparse_cache_entries) if sparse else 0
    if sparse and world > 1:
        ce = torch.tensor([cache_entries], device="cuda", dtype=torch.int64)
        dist.all_reduce(ce, op=dist.ReduceOp.MAX)
        cache_entries = int(ce[0])  # the largest rank bounds the step

    # ---- e2e leg: host buffers through the C ABI ----
    def e2e_once():
        set_data()
        h.generate_q()
        return h.solver_cg(b, None, QA, K, 0.0)

    if a.skip_e2e:
        t_e2e_ms, st2 = float("nan"), h.stats()
    else:
        e2e_once()  # warm-up (allocations)
        barrier()
        h.reset_stats()
        f0, f1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        f0.record(stream)
        e2e_once()
        f1.record(stream)
        barrier()
        t_e2e_ms = f0.elapsed_time(f1)
        st2 = h.stats()
    sampler.stop_flag.set()
    sampler.join(timeout=2)

    if world > 1:
        tt = torch.tensor([t_ms, t_e2e_ms], device="cuda", dtype=torch.float64)
        dist.all_reduce(tt, op=dist.ReduceOp.MAX)
        t_ms, t_e2e_ms = float(tt[0]), float(tt[1])

    if rank == 0:
        peaks, peak_kind = measured_peaks()
        flops = alg_flops_per_matvec(n, F) / world  # per launch on ONE GPU: the tile list is split evenly over the ranks
        achieved = flops / (matvec_ms * 1e-3) / 1e12 if matvec_ms > 0 else 0.0
        # the kernel is timed inside a multi-second solve -> sustained tensor figure
        peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                "peak_source": f"{peak_kind} bf16 dense (sustained)", "kernel": {1: "k_kmv_tile (SIMT fp32)", 2: "k_gram_tc (tcgen05 fp16x2-split)", 3: "gemv pair"}.get(path, "?"),
                "alg_flops_per_launch": flops, "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
        if not sparse and path == 2 and dense_cache_bytes > 0:
            # cached kernel matrix: the per-iteration kernel streams the stored lower-triangle tiles once (SURVEY.md §8d "cached dense K")
            hb = float(peaks.get("hbm_gbs", 6650.0))
            ach = dense_cache_bytes / (matvec_ms * 1e-3) / 1e9
            asm_tf = flops / (build_ms * 1e-3) / 1e12 if build_ms > 0 else None
            roof = {"bound": "hbm", "achieved": ach, "peak": hb, "unit": "GB/s", "frac": ach / hb, "traffic": None, "peak_source": f"{peak_kind} HBM copy",
                    "kernel": "k_cached_gemv (sweep over the stored kernel-matrix tiles; k_pack / k_tc_reduce inside the timed span)",
                    "alg_bytes_per_launch": float(dense_cache_bytes), "ms_per_launch": matvec_ms, "launches_timed": n_matvecs,
                    "assembly": {"kernel": "k_gram_tc<STORE> (tcgen05 fp16x2-split, once per data set, inside the timed region; it also serves the first matvec)", "ms": build_ms,
                                 "alg_flops": flops, "achieved_tflops": asm_tf, "frac_of_bf16_sustained": asm_tf / peak if asm_tf else None,
                                 "frac_issued": 2.0 * asm_tf / peak if asm_tf else None}}
            if world == 1 and (n, F, a.kernel, a.dtype) == (65536, 1024, "rbf", "f32"):
                # dram__bytes_read.sum + dram__bytes_write.sum of one k_cached_gemv launch of this exact workload (ncu --set full capture)
                roof["traffic"] = 8.625233e9 + 0.068885e9
                roof["traffic_source"] = "profiles/r1_cached_gemv_v2_ncu_full_raw.csv"
        elif not sparse and path == 2:
            # the fp16x2 split issues two pass-equivalents of fp16-rate MMA per algorithmic flop (hi.hi in fp16, both cross terms as one
            # fp8 chain of twice the K extent at twice the rate): `frac` stays on ALGORITHMIC flops, this is the tensor pipe's own view
            roof["issued_fp16_equiv_tflops"] = 2.0 * achieved
            roof["frac_issued"] = 2.0 * achieved / peak
        if not sparse and world == 1 and path == 2 and dense_cache_bytes == 0 and (n, F, a.kernel, a.dtype) == (65536, 1024, "rbf", "f32"):
            # dram__bytes_read.sum + dram__bytes_write.sum of one k_gram_tc launch of this exact workload (ncu --set full capture)
            roof["traffic"] = 2.807074e9 + 0.102859e9
            roof["traffic_source"] = "profiles/r1_gram_tc_final_ncu_full_raw.csv"
        if sparse and cache_entries > 0:
            # cached S: the matvec streams (value + 16-bit index) per stored entry, once (SURVEY.md §8d "cached S" row)
            nbytes = float(cache_entries) * (np.dtype(dt).itemsize + 2)
            ach = nbytes / (matvec_ms * 1e-3) / 1e9
            hb = float(peaks.get("hbm_gbs", 6650.0))
            roof = {"bound": "hbm", "achieved": ach, "peak": hb, "unit": "GB/s", "frac": ach / hb, "traffic": None, "peak_source": f"{peak_kind} HBM copy",
                    "kernel": "k_sp_cached_mv (SpMV over the cached S entries; weights / fold / finish kernels inside the timed span)",
                    "alg_bytes_per_launch": nbytes, "entries_per_gpu": cache_entries, "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
            flops = 2.0 * cache_entries
        elif sparse:
            roof = {"bound": "fp", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None,
                    "kernel": "k_sp_kmv (S recomputed on the fly: the cache did not fit or was disabled)", "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
            flops = 0.0
        line = {
            "metric": METRIC, "value": K / (t_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": t_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": a.dtype, "data": "synthetic", "config": workload_config(a),
            "matvec_gflops": world * flops / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else None,
            "clocks": sampler.summary(),
            "e2e": {"value": K / (t_e2e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": st2.h2d_bytes / K, "d2h_bytes_per_step": st2.d2h_bytes / K,
                    "ms_total": t_e2e_ms},
            "gpu_launches": launches, "warmup_steps_run": warm_steps,
            "collectives_per_step": (st.collectives / K) if world > 1 else 0, "peer_exchanges_per_step": (st.peer_exchanges / K) if world > 1 else 0,
            "cache_build_ms": build_ms,
            "roofline": roof,
        }
        if world == 1 and not a.no_cpu_baseline and not sparse:
            res = cpu_reference_iteration_rate(a, 1, 0, X, a.cpu_sample_points, budget_s=12.0)
            line["cpu_baseline"] = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
        if a.skip_e2e:
            line["e2e"] = None
        print(json.dumps(line), flush=True)
    h.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)


if __name__ == "__main__":
    main()
