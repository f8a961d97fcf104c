"""Accuracy / speed of the two tcgen05 cross-term modes against the f64 oracle (design input for DESIGN.md §3.1).
usage: python tools/tc_mode_numerics.py [points] [features]"""
import json
import sys

import numpy as np

sys.path.insert(0, ".")
from oracle import port  # noqa: E402
from plssvm_sparse_fp22_b200 import capi, datagen  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
F = int(sys.argv[2]) if len(sys.argv) > 2 else 1024
res = {}
for kernel in ("rbf", "polynomial"):
    for problem in ("blobs", "planes"):
        X, y = datagen.make_dense(n, F, seed=43, problem=problem, dtype=np.float32)
        d = np.random.default_rng(1).standard_normal(n - 1).astype(np.float32)
        X64 = X.astype(np.float64)
        q64 = port.generate_q(X64, kernel)
        qa = port.kernel_function(X64[-1], X64[-1], kernel, gamma=1.0 / F) + 1.0
        want = port.matvec(X64, q64, d.astype(np.float64), qa, 1.0, kernel, rows=(0, n - 1))
        for name, path, mode in (("simt_fp32", capi.PATH_SIMT, 0), ("tc_f16x3", capi.PATH_TC, capi.TC_F16X3), ("tc_fp8x", capi.PATH_TC, capi.TC_FP8X)):
            with capi.Handle(np.float32, kernel, gamma=1.0 / F) as h:
                h.set_option(capi.OPT_MATVEC_PATH, path)
                if mode:
                    h.set_option(capi.OPT_TC_MODE, mode)
                h.set_data(X)
                h.set_q(q64.astype(np.float32), qa)
                h.matvec(d)
                h.reset_stats()
                got = h.matvec(d)
                st = h.stats()
                a, rho, it = h.learn(y, eps=1e-3)
            a_o, rho_o, it_o = port.learn(X64, y.astype(np.float64), kernel, eps=1e-3, use_rows=True)
            res[f"{kernel}/{problem}/{name}"] = {
                "matvec_relerr": float(np.linalg.norm(got - want) / np.linalg.norm(want)), "ms": st.matvec_ms / max(st.matvecs_timed, 1),
                "iters": it, "iters_oracle": it_o, "alpha_relerr": float(np.linalg.norm(a - a_o) / np.linalg.norm(a_o))}
print(json.dumps(res, indent=1))
