"""Quick GPU check of the tcgen05 path (path 2) against the SIMT path (path 1) and the f64 oracle."""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from oracle import port  # noqa: E402
from plssvm_sparse_fp22_b200 import capi, datagen  # noqa: E402


def run(n, F, kernel, problem="blobs"):
    X, y = datagen.make_dense(n, F, seed=n + F, problem=problem, dtype=np.float32)
    rng = np.random.default_rng(1)
    d = rng.uniform(-1, 2, n - 1).astype(np.float32)
    res = {}
    for path in (1, 2):
        with capi.Handle(np.float32, kernel, 3, 1.0 / F, 0.5 if kernel == "polynomial" else 0.0, 1.0) as h:
            h.set_option(capi.OPT_MATVEC_PATH, path)
            h.set_data(X)
            q = h.generate_q()
            qa = h.kxx_last() + 1.0
            h.set_q(q, qa)
            t0 = time.perf_counter()
            out = h.matvec(d, 1.0)
            dt = time.perf_counter() - t0
            st = h.stats()
            res[path] = (out, st.matvec_path, st.matvec_ms, dt)
    X64 = X.astype(np.float64)
    q64 = port.generate_q(X64, kernel, 3, 1.0 / F, 0.5 if kernel == "polynomial" else 0.0)
    qa64 = port.kernel_function(X64[-1], X64[-1], kernel, 3, 1.0 / F, 0.5 if kernel == "polynomial" else 0.0) + 1.0
    want = port.matvec(X64, q64, d.astype(np.float64), qa64, 1.0, kernel, 3, 1.0 / F, 0.5 if kernel == "polynomial" else 0.0, 1.0, rows=(0, n - 1)) if n <= 3000 else None
    nr = lambda a, b: float(np.linalg.norm(a.astype(np.float64) - b.astype(np.float64)) / np.linalg.norm(b.astype(np.float64)))
    msg = f"n={n} F={F} {kernel}: paths={res[1][1]},{res[2][1]} simt_vs_tc={nr(res[2][0], res[1][0]):.2e} ms simt={res[1][2]:.3f} tc={res[2][2]:.3f}"
    if want is not None:
        msg += f" | vs f64 oracle: simt={nr(res[1][0], want):.2e} tc={nr(res[2][0], want):.2e}"
    print(msg, flush=True)


if __name__ == "__main__":
    for (n, F) in [(300, 64), (257, 33), (1000, 200), (2049, 1024), (700, 130)]:
        for k in ("rbf", "polynomial"):
            run(n, F, k)
    run(16384, 1024, "rbf")
    run(16384, 1024, "polynomial")
    run(2049, 200, "rbf", "planes")
