"""Evidence for the float tolerance policy: for every float case of tests/golden/cg_synth.npz print the reference's own
float result (1 thread) against the f64 truth, the reference's 1-vs-8-thread float spread, and this library's float result.
Run on a B200:  python tools/float_policy_table.py > profiles/r2_float_policy_table.md"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from plssvm_sparse_fp22_b200 import capi, datagen  # noqa: E402


def nrm_rel(a, b):
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    return float(np.linalg.norm(a - b) / max(np.linalg.norm(a), 1e-300))


g = np.load(os.path.join(ROOT, "tests", "golden", "cg_synth.npz"))
print("| case | kernel | eps | iters ref f32 (1 thr / 8 thr) | iters ours | ref f32 err vs f64 truth | ref f32 thread spread | ours err vs f64 truth | ours / max(ref err, spread) |")
print("|---|---|---|---|---|---|---|---|---|")
worst = 0.0
for case in ("blobs_257x33_s7", "planes_384x64_s8", "blobs_1000x256_s43"):
    problem, shape, seed = case.split("_")
    n, F = map(int, shape.split("x"))
    X, y = datagen.make_dense(n, F, seed=int(seed[1:]), problem=problem)
    for kernel in ("linear", "polynomial", "rbf"):
        for eps in (1e-3, 1e-6):
            key = f"{case}/f32/{kernel}/{eps:g}"
            truth = g[f"{case}/f64/{kernel}/{eps:g}/alpha"]
            for path in ((0,) if kernel == "linear" else (0, 1)):
                with capi.Handle(np.float32, kernel, gamma=1.0 / F) as h:
                    if path:
                        h.set_option(capi.OPT_MATVEC_PATH, path)
                    h.set_data(X.astype(np.float32))
                    alpha, rho, it = h.learn(y.astype(np.float32), eps=eps)
                    pth = h.stats().matvec_path
                err_ref = nrm_rel(truth, g[key + "/alpha"])
                spread = float(g[key + "/spread_alpha"])
                err = nrm_rel(truth, alpha)
                ratio = err / max(err_ref, spread, 1e-30)
                worst = max(worst, ratio) if err > 1e-5 else worst
                print(f"| {case} | {kernel} (path {pth}) | {eps:g} | {int(g[key + '/iters'])} / {int(g[key + '/iters8'])} | {it} | {err_ref:.2e} | {spread:.2e} | {err:.2e} | {ratio:.2f} |")
print(f"\nworst ratio among the cases above the 1e-5 floor: {worst:.2f}")
