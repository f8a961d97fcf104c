#!/usr/bin/env python
"""bench.py — CG iterations/sec of the LS-SVM hot path (BASELINE.json metric) on N B200s of one node.

Main line (every N): BASELINE configs[1] — RBF kernel, 65536 points x 1024 dense features, float, gamma = 1/F, C = 1 — in
fixed-iteration mode (eps = 0) because at the configured eps = 1e-3 CG converges after 1-2 iterations.  A "step" is
one CG iteration (implicit matvec Abar*d + the fused vector updates / dot products).  A timed solve of K steps contains K + 1
applications of the implicit matrix (the initial residual, csvm.cpp:94) plus one more per 50 iterations (csvm.cpp:130-135), and the
one-off assembly of the cached kernel-matrix tiles (dropped before every timed solve) — nothing is skipped.

  value : K / t, t = device time (CUDA events on the launching stream, max over ranks) of one b200svm_cg_solve call of exactly K
          steps with the training matrix already resident in HBM.  The solve is repeated R times (each bracketed by barrier +
          synchronize, cache dropped in between) until >= 0.5 s have been timed; t is the MEDIAN repetition (min / max reported).
  e2e   : the same through the C ABI with HOST buffers: b200svm_set_data_dense (H2D of X from pinned memory) +
          b200svm_generate_q + b200svm_cg_solve (H2D of b, D2H of alpha) inside the timed region.
  roofline : the dominant kernel of the step — the HBM-bound sweep over the stored kernel-matrix tiles (k_tile_sweep; algorithmic
          bytes = the stored tiles) or, with --dense-cache 1, the tcgen05 Gram-tile kernel (algorithmic flops F m (m+1)); duration
          from CUDA events recorded around each matvec inside libb200svm (b200svm_stats.matvec_ms / matvecs).
  parity (N > 1): alpha bit-identical on every rank, and its norm-wise distance to the alpha of a single-GPU solve of the same K steps.
  extra_workloads : the other BASELINE configs, each with its own roofline / e2e / clocks:
          N = 1   configs[2] geometry scaled to one GPU (131072 x 100k, 1e-3, rbf, double) and configs[3] (262144 x 20k, 1e-2,
                  polynomial, FP22 rows) when its S fits;  N = 2, 4  configs[3];  N = 8  configs[2] (1M x 100k), configs[3] and
                  configs[4] (predict, 4M points x 262144 support vectors).   --extra none switches them off.
  cpu_baseline / --impl reference : the UNMODIFIED reference OpenMP backend (oracle/_ref, built from /root/reference) timed on the
          host cores.  --impl reference runs ONE run_device_kernel call (= the matvec of one CG iteration) at the FULL configs[1]
          shape on the same X as the GPU arm when a probe predicts that it fits --ref-budget-s (else the largest row subset that
          does, scaled by the pair count and labelled so).
"""
from __future__ import annotations

import argparse
import json
import math
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
    ap.add_argument("--cpu-sample-points", type=int, default=0, help="row subset for the CPU baseline (0 = calibrate)")
    ap.add_argument("--ref-budget-s", type=float, default=900.0, help="--impl reference: CPU seconds the one timed call may take")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-cuda", action="store_true", help="skip the reference-CUDA-backend baseline leg (N = 1)")
    ap.add_argument("--seed", type=int, default=43)
    ap.add_argument("--density", type=float, default=0.0,
                    help="> 0: the MAIN line becomes a sparse workload (CSR rows of this density) instead of configs[1]")
    ap.add_argument("--storage", default="native", choices=["native", "fp22"], help="sparse value storage")
    ap.add_argument("--dense-cache", type=int, default=0,
                    help="dense tcgen05 path: 0 store the kernel-matrix tiles on first use if they fit (matvec = HBM-bound sweep), "
                         "1 recompute the Gram tiles on the tensor cores in every matvec")
    ap.add_argument("--peer-exchange", type=int, default=0, help="multi-GPU: 0 peer-memory exchange (default), 1 ncclAllReduce")
    ap.add_argument("--skip-e2e", action="store_true", help="skip the host-buffer leg (e2e = null)")
    ap.add_argument("--cache", type=int, default=0, help="sparse: 0 cache S if it fits, 1 recompute on the fly, 2 require the cache")
    ap.add_argument("--extra", default="auto", help="extra workloads: auto | none | comma list of cfg2,cfg2s,cfg3,cfg4")
    ap.add_argument("--min-timed-s", type=float, default=0.5, help="repeat the K-step solve until this much device time is timed")
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
        "dense_cache": "auto (kernel-matrix tiles stored once inside every timed solve)" if a.dense_cache == 0 else "off (Gram tiles recomputed every matvec)",
    }


def alg_flops_per_matvec(n, F):
    m = n - 1
    return float(F) * m * (m + 1)  # 2 F P, P = m(m+1)/2 unordered pairs incl. diagonal


# ----------------------------------------------------------------------------------------------------------------------
# CPU arm: the reference OpenMP backend (oracle/_ref) or the oracle port
# ----------------------------------------------------------------------------------------------------------------------
def cpu_matvec_seconds(X, ns, kernel, dt, use_ref):
    """One run_device_kernel call (ret += Abar d, svm_kernel.cpp:21-47) on the first ns points, all features."""
    from oracle import port, ref
    Xs = np.ascontiguousarray(X[:ns], dt)
    d = np.ones(ns - 1, dt)
    q = np.zeros(ns - 1, dt)
    t0 = time.perf_counter()
    if use_ref:
        ref.matvec(Xs, q, d, 1.0, 1.0, kernel)
    else:
        port.matvec(Xs, q, d, 1.0, 1.0, kernel, rows=(0, ns - 1))
    return time.perf_counter() - t0


def cpu_reference_rate(a, X, budget_s, sample_points=0, full_if_fits=False):
    """CG iterations/s of the reference on the host cores.  One iteration = one run_device_kernel call (the O(m) vector algebra is
    ignored: < 0.1 %).  A probe on 1024 points predicts the cost; the timed call uses the full shape when `full_if_fits` and the
    prediction fits the budget, else the largest row subset that does (scaled by the pair count, labelled 'extrapolated')."""
    from oracle import ref
    dt = np.float32 if a.dtype == "f32" else np.float64
    use_ref = ref.available()
    cores = os.cpu_count() or 1
    if use_ref:
        ref.set_num_threads(cores)
    n = X.shape[0]
    pairs = (lambda k: (k - 1) * k / 2.0) if use_ref else (lambda k: float(k - 1) * (k - 1))
    probe_n = min(n, 1024)
    cpu_matvec_seconds(X, probe_n, a.kernel, dt, use_ref)  # first touch / thread start-up
    t_probe = cpu_matvec_seconds(X, probe_n, a.kernel, dt, use_ref)
    per_pair = t_probe / max(pairs(probe_n), 1.0)
    predicted_full = per_pair * pairs(n)
    if sample_points > 0:
        ns = min(n, sample_points)
    elif full_if_fits and predicted_full <= budget_s:
        ns = n
    else:
        ns = int(min(n, max(probe_n, math.sqrt(2.0 * budget_s / per_pair) if use_ref else math.sqrt(budget_s / per_pair))))
        if ns < n:
            ns = max(probe_n, ns // 256 * 256)
    t = cpu_matvec_seconds(X, ns, a.kernel, dt, use_ref)
    t_full = t * pairs(n) / pairs(ns)
    full = ns == n
    return {
        "value": 1.0 / t_full, "unit": UNIT, "cores": cores, "kind": "reference" if use_ref else "port",
        "sample": (f"full shape: one run_device_kernel call on all {n} points x {a.features} features, measured ({t:.1f} s CPU wall)" if full else
                   f"run_device_kernel on the first {ns} of {n} points (all {a.features} features), one timed call ({t:.1f} s CPU wall), "
                   "extrapolated by the pair count to one full iteration"),
        "ms_per_step": 1e3 * t_full, "seconds_measured": t, "full_shape": full,
        "cross_check": {"points": probe_n, "seconds": t_probe, "predicted_full_shape_s": predicted_full},
    }


def run_reference_arm(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from plssvm_sparse_fp22_b200 import datagen
    dt = np.float32 if a.dtype == "f32" else np.float64
    X, _ = datagen.make_dense(a.points, a.features, seed=a.seed, problem="blobs", dtype=dt)  # the same X as the GPU arm
    t0 = time.perf_counter()
    res = cpu_reference_rate(a, X, a.ref_budget_s, a.cpu_sample_points, full_if_fits=True)
    line = {
        "metric": METRIC, "value": res["value"], "unit": UNIT, "n_gpus": a.gpus, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": res["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": a.dtype, "data": "synthetic", "config": workload_config(a), "impl": "reference",
        "cpu_baseline": {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")},
        "e2e": {"value": res["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "steps_measured": 1, "full_shape": res["full_shape"], "seconds_measured": res["seconds_measured"], "cross_check": res["cross_check"],
        "wall_s": time.perf_counter() - t0,
        "note": "one step = one run_device_kernel call of the unmodified reference (its CG iteration is this call plus O(m) vector "
                "algebra); --steps / --warmup are accepted but a single call is timed: it lasts minutes on the host cores",
    }
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------------------------------------------
# GPU arm
# ----------------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """SM clock and throttle reasons DURING the timed regions.  NVML in-process (a sample every ~5 ms); `nvidia-smi` (one sample
    per ~100 ms) if NVML cannot be initialised."""
    Q = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
    NAMES = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index = index
        self.samples = []   # (sm_mhz, sm_max_mhz, [active reason names])
        self.stop_flag = threading.Event()
        self.active = threading.Event()  # samples are kept only while a timed region is open
        self.source = "nvidia-smi"
        self.nv = None
        try:
            import pynvml
            pynvml.nvmlInit()
            vis = os.environ.get("CUDA_VISIBLE_DEVICES", "")
            phys = index
            tok = ""
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
            if not self.active.is_set():
                self.stop_flag.wait(0.002)
                continue
            try:
                if self.nv is not None:
                    sm = float(self.nv.nvmlDeviceGetClockInfo(self.dev, self.nv.NVML_CLOCK_SM))
                    try:
                        bits = int(self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.dev))
                    except Exception:
                        bits = int(self.nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.dev))
                    try:
                        mem = float(self.nv.nvmlDeviceGetClockInfo(self.dev, self.nv.NVML_CLOCK_MEM))
                        pw = float(self.nv.nvmlDeviceGetPowerUsage(self.dev)) / 1e3
                    except Exception:
                        mem, pw = None, None
                    self.samples.append((sm, self.sm_max, [nm for nm, m in self.masks if bits & m], mem, pw))
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

    def take(self):
        """Summary of the samples since the last call."""
        s, self.samples = self.samples, []
        sm = [x[0] for x in s]
        mx = [x[1] for x in s]
        reasons = [nm for nm in self.NAMES if any(nm in x[2] for x in s)]
        mem = [x[3] for x in s if len(x) > 3 and x[3] is not None]
        pw = [x[4] for x in s if len(x) > 4 and x[4] is not None]
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(s), "source": self.source, "sm_mhz_min": min(sm) if sm else None, "mem_mhz_min": min(mem) if mem else None,
                "mem_mhz": float(np.median(mem)) if mem else None, "power_w_median": float(np.median(pw)) if pw else None, "power_w_max": max(pw) if pw else None}


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return json.load(open(p)), "measured"
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0}, "fallback (B200_PROFILING.md)"


class Env:
    """Process-wide state of the GPU arm: ranks, stream, barrier, clock sampler, timing helpers."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py: no CUDA device — the B200 arm has no CPU fallback")
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))
        self.stream = torch.cuda.current_stream()
        self.sampler = ClockSampler(self.local)
        self.sampler.start()
        self.peaks, self.peak_kind = measured_peaks()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def fresh_uid(self, capi):
        torch = self.torch
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if self.rank == 0:
            idbuf.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
        self.dist.broadcast(idbuf, 0)
        return bytes(idbuf.cpu().numpy().tobytes())

    def max_over_ranks(self, values):
        if self.world == 1:
            return [float(v) for v in values]
        t = self.torch.tensor(list(values), device="cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return [float(v) for v in t.cpu()]

    def timed(self, fn, before=None):
        """Device time of fn() on the launching stream between two barriers; returns (ms on this rank, result)."""
        torch = self.torch
        if before is not None:
            before()
        self.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        self.sampler.active.set()
        e0.record(self.stream)
        out = fn()
        e1.record(self.stream)
        self.barrier()
        self.sampler.active.clear()
        return e0.elapsed_time(e1), out

    def repeat_timed(self, fn, before, min_total_s, rmin, rmax):
        """Repeat fn until min_total_s of device time (max over ranks) have been timed; the repeat count is agreed over the ranks.
        Returns (per-repetition ms as max over ranks, last result)."""
        t0, out = self.timed(fn, before)
        all_ts = [self.max_over_ranks([t0])[0]]
        # batches of repetitions until enough device time has been timed (the first repetition can be an outlier, so the count is
        # re-estimated from the median after every batch); every decision uses values that are identical on all ranks
        while len(all_ts) < rmax:
            total = sum(all_ts)
            if len(all_ts) >= rmin and total >= min_total_s * 1e3:
                break
            est = float(np.median(all_ts))
            more = int(min(rmax - len(all_ts), max(rmin - len(all_ts), math.ceil(max(min_total_s * 1e3 - total, 0.0) / max(est, 1e-3)), 1)))
            ts = []
            for _ in range(more):
                t, out = self.timed(fn, before)
                ts.append(t)
            all_ts += self.max_over_ranks(ts)
        return all_ts, out

    def close(self):
        self.sampler.stop_flag.set()
        self.sampler.join(timeout=2)
        if self.world > 1:
            self.dist.destroy_process_group()


def spread(ts, K):
    return {"reps": len(ts), "ms_per_step_median": float(np.median(ts)) / K, "ms_per_step_min": min(ts) / K, "ms_per_step_max": max(ts) / K,
            "timed_region_s": sum(ts) / 1e3}


def warm_up(env, solve, W):
    """Untimed: solves of max(W, 3) steps until the device has been busy for ~0.3 s.  The first solve carries one-off costs (IPC /
    NCCL set-up, allocations, cache assembly), so the repeat count is taken from the SECOND one and agreed over the ranks."""
    steps = max(W, 3)
    solve(steps)
    env.barrier()
    t0 = time.perf_counter()
    solve(steps)
    env.torch.cuda.synchronize()
    dt = env.max_over_ranks([time.perf_counter() - t0])[0]
    reps = int(min(60, max(2, 0.3 / max(dt, 1e-4))))
    for _ in range(reps):
        solve(steps)
    return (2 + reps) * steps


def ncu_traffic(csv_name, kernel_prefix):
    """dram__bytes_read.sum + dram__bytes_write.sum (bytes, one launch) of the first launch of `kernel_prefix` in a committed
    `ncu --set full ... --page raw --csv` export under profiles/; None when the file or the kernel is missing."""
    import csv
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    try:
        with open(os.path.join(ROOT, "profiles", csv_name), newline="") as f:
            rows = list(csv.reader(f))
        hdr, units = rows[0], rows[1]
        ik, ir, iw = hdr.index("Kernel Name"), hdr.index("dram__bytes_read.sum"), hdr.index("dram__bytes_write.sum")
        for r in rows[2:]:
            if kernel_prefix in r[ik]:
                return float(r[ir].replace(",", "")) * scale[units[ir]] + float(r[iw].replace(",", "")) * scale[units[iw]]
    except Exception:
        return None
    return None


def hbm_roofline(env, nbytes, matvec_ms, kernel, extra=None):
    hb = float(env.peaks.get("hbm_gbs", 6650.0))
    ach = nbytes / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else 0.0
    roof = {"bound": "hbm", "achieved": ach, "peak": hb, "unit": "GB/s", "frac": ach / hb, "traffic": None,
            "traffic_note": "not measured in this run (ncu captures of this kernel: profiles/)",
            "peak_source": f"{env.peak_kind} HBM copy", "kernel": kernel, "alg_bytes_per_launch": float(nbytes), "ms_per_launch": matvec_ms}
    if extra:
        roof.update(extra)
    return roof


def reference_cuda_baseline(a, X, b, K):
    """the UNMODIFIED reference CUDA backend (oracle/_ref/liboracle_ref_cuda.so, built for sm_100 from
    /root/reference) on the same GPU and the same data: one launch of its device_kernel_* (the matvec) timed with CUDA events, and
    its host-side CG (gpu_csvm::solver_CG, src/plssvm/backends/gpu_csvm.cpp:185-324) for a few fixed iterations."""
    from oracle import ref_cuda
    if not ref_cuda.available():
        return {"unavailable": "oracle/_ref/liboracle_ref_cuda.so missing (make -C oracle ref_cuda where /root/reference exists)"}
    try:
        m = X.shape[0] - 1
        q = np.zeros(m, X.dtype)
        d = np.ones(m, X.dtype)
        _, ms_mean, ms_best = ref_cuda.matvec_ms(X, q, d, 1.0, 1.0, a.kernel, reps=2)
        steps = min(K, 3)
        _, it, secs, setup = ref_cuda.solver_cg(X, b, q, 2.0, steps, 0.0, a.kernel)
        F = X.shape[1]
        return {"kind": "reference CUDA backend, unmodified, nvcc -arch=sm_100 -O3 --use_fast_math (its Release flags)",
                "kernel": {"linear": "device_kernel_linear", "polynomial": "device_kernel_poly", "rbf": "device_kernel_radial"}[a.kernel] +
                          " (src/plssvm/backends/CUDA/svm_kernel.cu)",
                "ms_per_matvec": ms_mean, "ms_per_matvec_best": ms_best, "alg_tflops": alg_flops_per_matvec(X.shape[0], F) / (ms_mean * 1e-3) / 1e12,
                "cg": {"value": steps / secs, "unit": UNIT, "steps": steps, "iterations_reported": it, "seconds": secs, "setup_data_on_device_s": setup,
                       "what": "gpu_csvm::solver_CG with eps = 0: steps + 1 kernel launches, CG vector algebra on the host, 4-5 PCIe copies per iteration"}}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"}


def run_main_workload(a, env):
    from plssvm_sparse_fp22_b200 import capi, datagen
    torch = env.torch
    rank, world = env.rank, env.world
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
        Xp = torch.empty((n, F), dtype=torch.float32 if dt == np.float32 else torch.float64, pin_memory=True)  # pinned host staging (e2e leg)
        Xp.numpy()[:] = X
        Xh = Xp.numpy()
    b = (y[:-1] - y[-1]).astype(dt)

    def set_data(hh):
        if not sparse:
            hh.set_data(Xh)
        elif a.storage == "fp22":
            hh.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
        else:
            hh.set_data_csr(rp, ci, vals, F)

    def drop_caches(hh):
        # derived data (cached kernel-matrix tiles / sparse S) is NOT carried into a timed region: its assembly is part of the job
        if sparse:
            hh.set_option(capi.OPT_SPARSE_CACHE, a.cache)
        else:
            hh.set_option(capi.OPT_DENSE_CACHE, a.dense_cache)

    def configure(hh):
        hh.set_stream(env.stream.cuda_stream)
        if a.path:
            hh.set_option(capi.OPT_MATVEC_PATH, a.path)
        drop_caches(hh)

    h = capi.Handle(dt, a.kernel, 3, 1.0 / F, 0.0, 1.0, device=env.local)
    configure(h)
    if world > 1:
        h.set_option(capi.OPT_PEER_EXCHANGE, a.peer_exchange)
        h.comm_init(env.fresh_uid(capi), rank, world)

    # ---- resident leg (value) ----
    set_data(h)
    h.generate_q()
    QA = h.kxx_last() + 1.0
    warm_steps = warm_up(env, lambda s: h.solver_cg(b, None, QA, s, 0.0), W)
    h.reset_stats()
    ts, (alpha, iters, delta) = env.repeat_timed(lambda: h.solver_cg(b, None, QA, K, 0.0), lambda: drop_caches(h), a.min_timed_s, 3, 200)
    st = h.stats()
    clocks = env.sampler.take()
    assert iters == K, (iters, K)
    R = len(ts)
    t_ms = float(np.median(ts))
    launches = int(st.kernel_launches)
    build_ms = float(st.cache_build_ms) / R
    dense_cache_bytes = int(st.dense_cache_bytes)
    # the matvec that assembles a cache is timed as a whole: take the assembly out to get the steady-state launch duration
    n_matvecs = int(st.matvecs_timed)
    if not sparse and dense_cache_bytes > 0 and build_ms > 0 and int(st.matvec_path) == 2:
        n_matvecs -= R  # dense tiles: the STORE pass of the tcgen05 kernel also serves the matvec that triggered it (no sweep in that one)
    matvec_ms = (st.matvec_ms - st.cache_build_ms) / max(n_matvecs, 1)
    matvec_total_ms = st.matvec_ms / R  # per solve, assembly included
    path = int(st.matvec_path)
    cache_entries = int(st.sparse_cache_entries) if sparse else 0
    if sparse and world > 1:
        cache_entries = int(env.max_over_ranks([cache_entries])[0])  # the largest rank bounds the step

    # ---- parity (N > 1): the configured solve (eps = 1e-3, imax = F) on N ranks and on one: bit-identical alpha on every rank, same
    #      iteration count, norm-wise distance.  (The K fixed iterations of the timed solve run far past convergence — delta / delta0
    #      reaches the float noise floor after 2-3 iterations — so their alpha is rounding noise and is not compared.)
    parity = None
    if world > 1:
        drop_caches(h)  # like the fresh single-GPU handle below, the solve starts with the assembly (STORE) pass
        aN, itN, _ = h.solver_cg(b, None, QA, F, 1e-3)
        t_alpha = torch.from_numpy(aN.astype(np.float64)).cuda()
        ref_alpha = t_alpha.clone()
        env.dist.broadcast(ref_alpha, 0)
        same = torch.tensor([1 if torch.equal(t_alpha, ref_alpha) else 0], device="cuda")
        env.dist.all_reduce(same, op=env.dist.ReduceOp.MIN)
        t_k = torch.from_numpy(alpha.astype(np.float64)).cuda()
        ref_k = t_k.clone()
        env.dist.broadcast(ref_k, 0)
        same_k = torch.tensor([1 if torch.equal(t_k, ref_k) else 0], device="cuda")
        env.dist.all_reduce(same_k, op=env.dist.ReduceOp.MIN)
        dist1, it1 = None, None
        if rank == 0:
            with capi.Handle(dt, a.kernel, 3, 1.0 / F, 0.0, 1.0, device=env.local) as h1:
                configure(h1)
                set_data(h1)
                h1.generate_q()
                a1, it1, _ = h1.solver_cg(b, None, h1.kxx_last() + 1.0, F, 1e-3)
            dist1 = float(np.linalg.norm(aN.astype(np.float64) - a1.astype(np.float64)) / np.linalg.norm(a1.astype(np.float64)))
        env.barrier()
        parity = {"alpha_bit_identical_across_ranks": bool(int(same[0]) == 1), "alpha_of_the_timed_solve_bit_identical_across_ranks": bool(int(same_k[0]) == 1),
                  "iterations": int(itN), "iterations_single_gpu": it1, "alpha_relerr_vs_single_gpu": dist1,
                  "note": "solve at the configured eps = 1e-3 (imax = F) on N ranks vs one rank: norm-wise ||alpha_N - alpha_1|| / ||alpha_1|| "
                          "(0 = bit-identical: the tile paths sum their partial vectors over eight canonical tile groups in a fixed tree for 1, 2, 4 and 8 GPUs)"}

    # ---- beside the headline: the same K-step solve with the Gram tiles recomputed in every matvec, and the CONFIGURED solve (eps = 1e-3:
    #      CG stops after 1-2 iterations) in both modes — the cached matrix only pays from the second matvec on ----
    side = None
    if not sparse and int(st.matvec_path) == 2 and a.dense_cache == 0:
        def solve_eps(mode):
            h.set_option(capi.OPT_DENSE_CACHE, mode)
            return h.solver_cg(b, None, QA, F, 1e-3)
        tc_ms, (_, it_c, _) = env.repeat_timed(lambda: solve_eps(capi.CACHE_AUTO), None, 0.0, 3, 3)
        tr_ms, (_, it_r, _) = env.repeat_timed(lambda: solve_eps(capi.CACHE_OFF), None, 0.0, 3, 3)
        tk_ms, _ = env.repeat_timed(lambda: (h.set_option(capi.OPT_DENSE_CACHE, capi.CACHE_OFF), h.solver_cg(b, None, QA, K, 0.0))[1], None, 0.0, 3, 3)
        h.set_option(capi.OPT_DENSE_CACHE, a.dense_cache)
        side = {"recompute_mode": {"value": K / (float(np.median(tk_ms)) * 1e-3), "unit": UNIT, "ms_per_step": float(np.median(tk_ms)) / K,
                                   "what": "the same K-step solve with B200SVM_OPT_DENSE_CACHE = 1: Gram tiles recomputed on the tensor cores in every matvec"},
                "configured_eps_1e-3": {"iterations": int(it_c), "iterations_recompute_mode": int(it_r), "ms_cached_mode": float(np.median(tc_ms)),
                                        "ms_recompute_mode": float(np.median(tr_ms)),
                                        "what": "one b200svm_cg_solve at the configured eps = 1e-3 (imax = F), data resident, assembly inside: with 1-2 iterations the "
                                                "cached mode is one STORE pass (= one recompute matvec) plus sweeps"}}

    # ---- e2e leg: host buffers through the C ABI ----
    def e2e_once():
        set_data(h)
        h.generate_q()
        return h.solver_cg(b, None, QA, K, 0.0)

    if a.skip_e2e:
        te, st2, clocks_e2e = None, h.stats(), None
    else:
        e2e_once()  # warm-up (allocations)
        h.reset_stats()
        te, _ = env.repeat_timed(e2e_once, None, a.min_timed_s, 3, 40)
        st2 = h.stats()
        clocks_e2e = env.sampler.take()

    fma_peak = h.probe_fma_tflops() if (not sparse and int(st.matvec_path) == 1) else None
    line = None
    if rank == 0:
        peaks = env.peaks
        flops = alg_flops_per_matvec(n, F) / world  # per launch on ONE GPU: the tile list is split evenly over the ranks
        achieved = flops / (matvec_ms * 1e-3) / 1e12 if matvec_ms > 0 else 0.0
        peak = float(peaks.get("bf16_tflops_sustained", peaks.get("bf16_tflops")))  # timed inside a long step -> sustained figure
        roof = {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak, "traffic": None,
                "peak_source": f"{env.peak_kind} bf16 dense (sustained)", "kernel": {1: "k_kmv_tile (SIMT)", 2: "k_gram_tc (tcgen05 fp16x2-split)", 3: "gemv pair"}.get(path, "?"),
                "alg_flops_per_launch": flops, "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
        matvec_gflops = world * flops / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else None
        if not sparse and path == 1 and dense_cache_bytes > 0:
            # dense double polynomial / rbf with the kernel matrix cached as double lower-triangle tiles
            fl_asm = float(F) * float(n - 1) * float(n) / world  # the STORE pass evaluates the lower triangle once: F m (m + 1) flops
            roof = hbm_roofline(env, dense_cache_bytes, matvec_ms, "k_tile_sweep_f64 (sweep over the stored double kernel-matrix tiles; k_sweep_reduce inside the timed span)",
                                {"launches_timed": n_matvecs,
                                 "assembly": {"kernel": "k_kmv_tile<double, STORE> (SIMT Gram tiles on the FP64 FMA pipe, once per data set, inside the timed region)",
                                              "ms": build_ms, "alg_flops": fl_asm, "achieved_tflops": fl_asm / (build_ms * 1e-3) / 1e12 if build_ms > 0 else None,
                                              "fp64_fma_peak_tflops_measured": fma_peak}})
            matvec_gflops = world * float(dense_cache_bytes) / 8.0 * 4.0 / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else None
        elif not sparse and path == 1:
            # SIMT Gram tiles (double precision, or float with --path 1): full square, no symmetry -> 2 F m^2 FMA flops per matvec,
            # bound by the FP64 / FP32 FMA pipe; the peak is measured on this device by a register-only FMA kernel
            fl = 2.0 * F * float(n - 1) * float(n - 1) / world
            ach = fl / (matvec_ms * 1e-3) / 1e12 if matvec_ms > 0 else 0.0
            roof = {"bound": "fp64-fma" if a.dtype == "f64" else "fp32-fma", "achieved": ach, "peak": fma_peak, "unit": "TFLOP/s", "frac": ach / fma_peak if fma_peak else None,
                    "traffic": None, "peak_source": "measured on this device: b200svm_probe_fma_tflops (16 independent FMA chains per thread, 2 flops per FMA)",
                    "kernel": "k_kmv_tile (SIMT register-tiled Gram tiles, full square)", "alg_flops_per_launch": fl,
                    "alg_flops_symmetric_minimum": flops, "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
            matvec_gflops = world * fl / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else None
        if not sparse and path == 2 and dense_cache_bytes > 0:
            # cached kernel matrix: the per-iteration kernel streams the stored lower-triangle tiles once (SURVEY.md §8d "cached dense K")
            asm_tf = flops / (build_ms * 1e-3) / 1e12 if build_ms > 0 else None
            roof = hbm_roofline(env, dense_cache_bytes, matvec_ms,
                                "k_tile_sweep (sweep over the stored kernel-matrix tiles; k_sweep_reduce (+ peer push) inside the timed span)",
                                {"launches_timed": n_matvecs,
                                 "assembly": {"kernel": "k_gram_tc<STORE> (tcgen05 fp16x2-split, once per data set, inside the timed region; it also serves the first matvec)",
                                              "ms": build_ms, "alg_flops": flops, "achieved_tflops": asm_tf, "frac_of_bf16_sustained": asm_tf / peak if asm_tf else None,
                                              "frac_issued": 2.0 * asm_tf / peak if asm_tf else None}})
            # DRAM traffic of one k_tile_sweep launch: from the committed ncu capture of this kernel when the run has the shape the
            # capture was taken on (configs[1] on one GPU: the same tiles, the same bytes); null otherwise — never a typed-in literal
            if world == 1 and a.points == 65536 and a.features == 1024 and a.kernel == "rbf" and a.dtype == "f32":
                tr = ncu_traffic("r2_tile_sweep_ncu_full_raw.csv", "k_tile_sweep")
                if tr is not None:
                    roof["traffic"] = tr
                    roof["traffic_over_algorithmic"] = tr / float(dense_cache_bytes)
                    roof["traffic_note"] = ("bytes per launch = dram__bytes_read.sum + dram__bytes_write.sum of k_tile_sweep in profiles/r2_tile_sweep_ncu_full_raw.csv "
                                            "(ncu --set full of this command at this shape, read at run time; not re-measured by this run)")
            # the sweep multiplies every stored entry twice (row part and mirrored column part): 4 flops per stored entry — NOT the
            # F m (m+1) of a recompute
            matvec_gflops = world * float(dense_cache_bytes) / (matvec_ms * 1e-3) / 1e9 if matvec_ms > 0 else None
        elif not sparse and path == 2:
            # the fp16x2 split issues two pass-equivalents of fp16-rate MMA per algorithmic flop: `frac` stays on ALGORITHMIC flops
            roof["issued_fp16_equiv_tflops"] = 2.0 * achieved
            roof["frac_issued"] = 2.0 * achieved / peak
        if sparse and cache_entries > 0:
            nbytes = float(cache_entries) * (np.dtype(dt).itemsize + 2)
            roof = hbm_roofline(env, nbytes, matvec_ms, "k_sp_cached_mv (SpMV over the cached S entries; weights / fold / finish kernels inside the timed span)",
                                {"entries_per_gpu": cache_entries, "launches_timed": n_matvecs})
            matvec_gflops = world * 2.0 * cache_entries / (matvec_ms * 1e-3) / 1e9
        elif sparse:
            roof = {"bound": "fp", "achieved": None, "peak": None, "unit": "TFLOP/s", "frac": None, "traffic": None,
                    "kernel": "k_sp_kmv (S recomputed on the fly: the cache did not fit or was disabled)", "ms_per_launch": matvec_ms, "launches_timed": n_matvecs}
            matvec_gflops = None
        line = {
            "metric": METRIC, "value": K / (t_ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": K, "warmup": max(W, 3),
            "ms_per_step": t_ms / K, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": a.dtype, "data": "synthetic", "config": workload_config(a),
            "timing": dict(spread(ts, K), rule="value = K / median over reps of the K-step solve (each rep: barrier + sync, events on the launching stream, max over ranks)"),
            "matvec_gflops": matvec_gflops,
            "clocks": clocks,
            "e2e": None if te is None else {"value": K / (float(np.median(te)) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": st2.h2d_bytes / len(te) / K,
                                            "d2h_bytes_per_step": st2.d2h_bytes / len(te) / K, "ms_total": float(np.median(te)),
                                            "timing": spread(te, K), "clocks": clocks_e2e},
            "gpu_launches": launches, "gpu_launches_per_step": launches / R / K, "warmup_steps_run": warm_steps,
            "collectives_per_step": (st.collectives / R / K) if world > 1 else 0, "peer_exchanges_per_step": (st.peer_exchanges / R / K) if world > 1 else 0,
            "cache_build_ms": build_ms,
            "breakdown_ms_per_solve": {"total": t_ms, "matvec_spans": matvec_total_ms, "of_which_cache_assembly": build_ms,
                                       "cg_vector_kernels": st.cg_vector_ms / R, "cg_vector_kernel_ms_each": st.cg_vector_ms / max(int(st.cg_vector_timed), 1),
                                       "launch_gaps_fill_copies": t_ms - matvec_total_ms - st.cg_vector_ms / R,
                                       "note": "rank 0, CUDA events inside libb200svm around every matvec (sweep + partial-vector reduction / peer push); "
                                               "cg_vector_kernels = the fused CG step kernel (several ranks: incl. waiting for the peers' slots); the rest is launch gaps and the initial fill / copies"},
            "roofline": roof,
        }
        if parity is not None:
            line["parity"] = parity
        if side is not None:
            line.update(side)
        if world == 1 and not a.no_cpu_baseline and not sparse:
            res = cpu_reference_rate(a, X, 12.0, a.cpu_sample_points)
            line["cpu_baseline"] = {k: res[k] for k in ("value", "unit", "cores", "kind", "sample")}
    if world == 1 and not a.no_ref_cuda and not sparse:
        h.close()  # free the HBM first
        ref_cuda = reference_cuda_baseline(a, X, b, K)
        if line is not None:
            line["baseline_ref_cuda"] = ref_cuda
    h.close()
    return line


# ----------------------------------------------------------------------------------------------------------------------
# extra workloads: BASELINE configs[2], [3] (sparse CG) and [4] (predict)
# ----------------------------------------------------------------------------------------------------------------------
def run_sparse_cg(env, tag, n, F, density, kernel, dtype, storage, K, seed, min_timed_s):
    from plssvm_sparse_fp22_b200 import capi, datagen
    rank, world = env.rank, env.world
    dt = np.float32 if dtype == "f32" else np.float64
    t0 = time.perf_counter()
    rp, ci, vals, y = datagen.make_sparse_csr(n, F, density, seed=seed, dtype=dt)
    fmt = None
    if storage == "fp22":
        vals = datagen.fp22_round(vals)
        rp_s, vals_s = datagen.pack_fp22_csr(rp, ci, vals)
        ci_s, fmt = None, capi.VAL_FP22
    else:
        rp_s, ci_s, vals_s = rp, ci, vals
    t_gen = time.perf_counter() - t0
    b = (y[:-1] - y[-1]).astype(dt)
    out = {"name": tag, "config": {"workload": f"{tag}: {kernel} kernel, {n} points x {F} features at density {density:g} (CSR, "
                                               f"{'FP22-packed rows, f32 compute' if storage == 'fp22' else dtype}), gamma=1/F, C=1, {world} GPU(s), rows / tiles of "
                                               "the implicit matrix partitioned over the ranks; fixed-iteration CG (eps=0) for the rate, eps=1e-3 for the learn() line",
                                   "points": n, "features": F, "density": density, "kernel": kernel, "storage": storage, "nnz": int(rp[-1])},
           "metric": METRIC, "unit": UNIT, "n_gpus": world, "steps": K, "dtype": dtype, "data": "synthetic", "datagen_s": t_gen}
    h = None
    try:
        h = capi.Handle(dt, kernel, 3, 1.0 / F, 0.0, 1.0, device=env.local)
        h.set_stream(env.stream.cuda_stream)
        if world > 1:
            h.comm_init(env.fresh_uid(capi), rank, world)

        def set_data():
            h.set_data_csr(rp_s, ci_s, vals_s, F, value_format=fmt)

        def drop():
            h.set_option(capi.OPT_SPARSE_CACHE, capi.CACHE_AUTO)

        set_data()
        h.generate_q()
        QA = h.kxx_last() + 1.0
        h.solver_cg(b, None, QA, 3, 0.0)  # warm-up: set-up of the exchange, allocations, one assembly
        h.solver_cg(b, None, QA, 3, 0.0)
        h.reset_stats()
        # value: data resident; S assembly + K steps inside the timed region
        ts, (alpha, iters, _) = env.repeat_timed(lambda: h.solver_cg(b, None, QA, K, 0.0), drop, min_timed_s, 2, 10)
        st = h.stats()
        clocks = env.sampler.take()
        R = len(ts)
        t_ms = float(np.median(ts))
        build_ms = float(st.cache_build_ms) / R
        n_mv = int(st.matvecs_timed)
        # (the matvec that triggers the assembly also sweeps: it stays in the count, its assembly time is taken out)
        matvec_ms = (st.matvec_ms - st.cache_build_ms) / max(n_mv, 1)
        entries = int(env.max_over_ranks([int(st.sparse_cache_entries)])[0])
        dense_bytes = int(env.max_over_ranks([int(st.dense_cache_bytes)])[0])

        # e2e: set_data_csr + generate_q + K steps from host buffers (assembly inside)
        def e2e_once():
            set_data()
            h.generate_q()
            return h.solver_cg(b, None, QA, K, 0.0)
        h.reset_stats()
        te, _ = env.repeat_timed(e2e_once, None, 0.0, 3, 3)  # median of three: one slow repetition (host hiccup) does not move it
        st2 = h.stats()
        # whole learn() at eps = 1e-3 from host buffers: set_data + q + CG to convergence + bias (wall clock, max over ranks)
        env.barrier()
        w0 = time.perf_counter()
        set_data()
        alpha_l, rho_l, it_l = h.learn(y, eps=1e-3)
        env.torch.cuda.synchronize()
        learn_s = env.max_over_ranks([time.perf_counter() - w0])[0]
        same = None
        if world > 1:
            torch = env.torch
            ta = torch.from_numpy(alpha_l.astype(np.float64)).cuda()
            ra = ta.clone()
            env.dist.broadcast(ra, 0)
            sm = torch.tensor([1 if torch.equal(ta, ra) else 0], device="cuda")
            env.dist.all_reduce(sm, op=env.dist.ReduceOp.MIN)
            same = bool(int(sm[0]) == 1)
        if dense_bytes > 0:
            roof = hbm_roofline(env, dense_bytes, matvec_ms, "k_tile_sweep (S is densely filled: stored as dense lower-triangle tiles, no indices)",
                                {"dense_tile_bytes_per_gpu": dense_bytes})
        elif entries > 0:
            roof = hbm_roofline(env, float(entries) * (np.dtype(dt).itemsize + 2), matvec_ms,
                                "k_sp_cached_mv (SpMV over the cached S entries: value + 16-bit index per entry)", {"entries_per_gpu": entries})
        else:
            roof = {"bound": "fp", "achieved": None, "peak": None, "unit": None, "frac": None, "traffic": None,
                    "kernel": "k_sp_kmv (S recomputed on the fly: the cache did not fit)", "ms_per_launch": matvec_ms}
        out.update({"value": K / (t_ms * 1e-3), "ms_per_step": t_ms / K, "timing": spread(ts, K), "cache_build_ms": build_ms, "ms_per_matvec": matvec_ms,
                    "roofline": roof, "clocks": clocks,
                    "e2e": {"value": K / (float(np.median(te)) * 1e-3), "unit": UNIT, "h2d_bytes_per_step": st2.h2d_bytes / len(te) / K,
                            "d2h_bytes_per_step": st2.d2h_bytes / len(te) / K, "ms_total": float(np.median(te)), "timing": spread(te, K)},
                    "learn": {"eps": 1e-3, "iterations": int(it_l), "wall_s": learn_s, "rho": float(rho_l), "alpha_bit_identical_across_ranks": same,
                              "what": "b200svm_set_data_csr + b200svm_learn from host buffers (q, S assembly, CG to delta <= eps^2 delta0, bias), wall clock, max over ranks"},
                    "gpu_launches": int(st.kernel_launches)})
    except Exception as e:  # an extra workload must not take the main line down
        out["error"] = f"{type(e).__name__}: {e}"
    if h is not None:
        h.close()
    return out


def run_predict(env, tag, svs, F, density, points, seed):
    from plssvm_sparse_fp22_b200 import capi, datagen
    rank, world = env.rank, env.world
    dt = np.float32
    t0 = time.perf_counter()
    rp, ci, v, _ = datagen.make_sparse_csr(svs, F, density, seed=seed, dtype=dt)
    share = points // world
    prp, pci, pv, _ = datagen.make_sparse_csr(share, F, density, seed=seed + 1 + rank, dtype=dt, min_nnz=0)
    t_gen = time.perf_counter() - t0
    rng = np.random.default_rng(seed)
    alpha = (rng.standard_normal(svs) * 0.1).astype(dt)  # alpha ~ N(0, 1e-2), rho = 0.1 (SURVEY.md §8d cfg 5)
    out = {"name": tag, "metric": "predict_points_per_sec", "unit": "points/s", "n_gpus": world, "dtype": "f32", "data": "synthetic", "datagen_s": t_gen,
           "config": {"workload": f"{tag}: predict, {share * world} sparse test points x {svs} support vectors, {F} features at density {density:g}, rbf, f32; "
                                  "test rows split over the ranks (every rank draws its own share), no collective on the data path",
                      "points": share * world, "svs": svs, "features": F, "density": density}}
    try:
        with capi.Handle(dt, "rbf", 3, 1.0 / F, 0.0, 1.0, device=env.local) as h:
            h.set_data_csr(rp, ci, v, F)
            w = min(share, 1024)
            h.predict_csr(alpha, 0.1, prp[: w + 1], pci[: prp[w]], pv[: prp[w]])  # warm-up
            h.reset_stats()
            env.barrier()
            env.sampler.active.set()
            w0 = time.perf_counter()
            dec = h.predict_csr(alpha, 0.1, prp, pci, pv)
            t = time.perf_counter() - w0
            env.barrier()
            env.sampler.active.clear()
            st = h.stats()
            atoms_gops = h.probe_smem_atomic_gops()
            fma_tflops = h.probe_fma_tflops(np.float32)
        t = env.max_over_ranks([t])[0]
        clocks = env.sampler.take()
        total = share * world
        evals = float(total) * svs
        products = evals * F * density * density  # expected sparse Gram products (P n F rho^2)
        # roofline of the on-the-fly kernel: one 32-bit shared-memory atomic add per Gram product (fixed-point accumulation, DESIGN.md
        # §4.1); peak = the rate measured on this device by b200svm_probe_smem_atomic_gops, times N GPUs
        peak = world * atoms_gops
        out.update({"value": total / t, "seconds": t, "kernel_evals_per_s": evals / t, "gram_products_per_s": products / t,
                    "roofline": {"bound": "smem-atomic", "achieved": products / t / 1e9, "peak": peak, "unit": "G atomic adds/s",
                                 "frac": products / t / 1e9 / peak if peak else None, "traffic": None,
                                 "kernel": "k_sp_kmv (S on the fly: one multiply + one shared-memory atomic add per Gram product, one ex2 per kernel value)",
                                 "peak_source": "measured on this device: b200svm_probe_smem_atomic_gops (ATOMS.ADD to pseudo-random slots of an 8192-entry accumulator) x N GPUs; "
                                                "algorithmic ops = expected Gram products P n F rho^2",
                                 "note": "timed through b200svm_predict_csr with host buffers: H2D of the points and D2H of the decision values are inside",
                                 "frac_of_fp32_fma_peak": (2.0 * products + 8.0 * evals) / t / 1e12 / (world * fma_tflops) if fma_tflops else None,
                                 "fp32_fma_peak_tflops_measured": fma_tflops,
                                 "fp32_note": "2 flops per Gram product + ~8 per kernel evaluation (exponent, ex2, weight) against the measured FP32 FMA peak x N"},
                    "clocks": clocks,
                    "e2e": {"value": total / t, "unit": "points/s", "h2d_bytes_per_step": int(st.h2d_bytes), "d2h_bytes_per_step": int(st.d2h_bytes)},
                    "gpu_launches": int(st.kernel_launches), "positive_labels_rank0": int((dec > 0).sum())})
    except Exception as e:
        out["error"] = f"{type(e).__name__}: {e}"
    return out


def extra_list(a, world):
    if a.extra == "none" or a.density > 0 or (a.points, a.features) != (65536, 1024):
        return []
    if a.extra != "auto":
        return [t.strip() for t in a.extra.split(",") if t.strip()]
    return {1: ["cfg2s", "cfg3"], 2: ["cfg3"], 4: ["cfg3"], 8: ["cfg2", "cfg3", "cfg4"]}.get(world, [])


def run_b200_arm(a):
    env = Env()
    line = run_main_workload(a, env)
    extras = []
    for tag in extra_list(a, env.world):
        if tag == "cfg2":
            extras.append(run_sparse_cg(env, "configs[2]", 1000000, 100000, 1e-3, "rbf", "f64", "native", 20, 44, a.min_timed_s))
        elif tag == "cfg2s":
            extras.append(run_sparse_cg(env, "configs[2] geometry at one-GPU size (1/8 of the rows)", 131072, 100000, 1e-3, "rbf", "f64", "native", 20, 44, a.min_timed_s))
        elif tag == "cfg3":
            extras.append(run_sparse_cg(env, "configs[3]", 262144, 20000, 1e-2, "polynomial", "f32", "fp22", 20, 45, a.min_timed_s))
        elif tag == "cfg4":
            extras.append(run_predict(env, "configs[4]", 262144, 20000, 1e-2, 4194304, 46))
    if env.rank == 0:
        line["extra_workloads"] = extras
        print(json.dumps(line), flush=True)
    env.close()


def main():
    a = parse_args()
    if a.impl == "reference":
        run_reference_arm(a)
    else:
        run_b200_arm(a)


if __name__ == "__main__":
    main()
