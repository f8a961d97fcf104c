"""Per-solve timing of the dense cached mode at configs[1] (65536 x 1024 f32 rbf): for a sequence of CG solves of K steps, with the
cache dropped or kept before each, print the assembly time and the mean sweep time reported by b200svm_stats, and the wall time of
the call.  Usage: python tools/probe_sweep.py [K]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from plssvm_sparse_fp22_b200 import capi, datagen  # noqa: E402

K = int(sys.argv[1]) if len(sys.argv) > 1 else 20
n, F = 65536, 1024
X, y = datagen.make_dense(n, F, seed=43, problem="blobs", dtype=np.float32)
b = (y[:-1] - y[-1]).astype(np.float32)
with capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0) as h:
    h.set_data(X)
    h.generate_q()
    qa = h.kxx_last() + 1.0
    plan = [("drop", 3), ("drop", K), ("keep", K), ("keep", K), ("drop", K), ("drop", K), ("set_data", K), ("set_data", K), ("keep", 3 * K)]
    for what, steps in plan:
        if what == "drop":
            h.set_option(capi.OPT_DENSE_CACHE, capi.CACHE_AUTO)
        elif what == "set_data":
            h.set_data(X)
            h.generate_q()
        h.reset_stats()
        t0 = time.perf_counter()
        h.solver_cg(b, None, qa, steps, 0.0)
        wall = (time.perf_counter() - t0) * 1e3
        st = h.stats()
        built = st.cache_build_ms > 0
        sweeps = st.matvecs_timed - (1 if built else 0)
        print(f"{what:9s} steps={steps:3d} wall={wall:7.2f} ms  build={st.cache_build_ms:6.2f} ms  sweeps={sweeps:3d}  "
              f"ms/sweep={(st.matvec_ms - st.cache_build_ms) / max(sweeps, 1):.3f}  launches={st.kernel_launches}", flush=True)
