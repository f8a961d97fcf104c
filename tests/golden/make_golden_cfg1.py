"""Golden vector for BASELINE configs[1] at FULL size from the unmodified reference (oracle/_ref): one CG iteration of
openmp::csvm::solver_CG (src/plssvm/backends/OpenMP/csvm.cpp:82-170) on the 65536 x 1024 float blobs data of bench.py — i.e. two
applications of the implicit matrix (initial residual + the first search direction).  ~10-15 minutes on 8 cores.

    python tests/golden/make_golden_cfg1.py            # writes tests/golden/cg_cfg1_full.npz
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref  # noqa: E402
from plssvm_sparse_fp22_b200 import datagen  # noqa: E402

n, F = 65536, 1024
X, y = datagen.make_dense(n, F, seed=43, problem="blobs", dtype=np.float32)
b = (y[:-1] - y[-1]).astype(np.float32)
ref.set_num_threads(os.cpu_count() or 1)
t0 = time.time()
q = ref.generate_q(X, "rbf")
kxx = ref.kernel_function(X[-1], X[-1], "rbf", gamma=1.0 / F)
QA = float(kxx) + 1.0
alpha, it, secs = ref.solver_cg(X, b, q, QA, 1, 1e-3, "rbf")
print(f"reference solver_CG: imax=1 -> iterations={it}, {secs:.1f} s, total {time.time() - t0:.1f} s, threads={ref.num_threads()}")
alpha64 = None
if "--f64" in sys.argv:
    # the same iteration in double (the "truth" the float policy of oracle/policy.py measures both float results against)
    X64 = X.astype(np.float64)
    q64 = ref.generate_q(X64, "rbf")
    QA64 = float(ref.kernel_function(X64[-1], X64[-1], "rbf", gamma=1.0 / F)) + 1.0
    alpha64, it64, secs64 = ref.solver_cg(X64, b.astype(np.float64), q64, QA64, 1, 1e-3, "rbf")
    print(f"reference solver_CG (double): iterations={it64}, {secs64:.1f} s")
extra = {} if alpha64 is None else {"alpha_f64": alpha64.astype(np.float64)}
np.savez_compressed(os.path.join(ROOT, "tests", "golden", "cg_cfg1_full.npz"), alpha=alpha, **extra, q_head=q[:4096], q_sum=np.float64(q.astype(np.float64).sum()),
                    QA_cost=np.float64(QA), iterations=np.int64(it), seconds=np.float64(secs), threads=np.int64(ref.num_threads()),
                    seed=np.int64(43), n=np.int64(n), F=np.int64(F))
