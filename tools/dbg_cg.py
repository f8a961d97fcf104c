import sys
import numpy as np
sys.path.insert(0, ".")
from oracle import port
from plssvm_sparse_fp22_b200 import capi, datagen
n, F = 300, 20
kernel = "polynomial"
dt = np.float32
X, y = datagen.make_dense(n, F, seed=77, problem="planes", dtype=dt)
q = port.generate_q(X, kernel)
qa = port.kernel_function(X[-1], X[-1], kernel, gamma=1.0 / F) + 1.0
b = (y[:-1] - y[-1]).astype(dt)
for path in (1, 2):
    with capi.Handle(dt, kernel, gamma=1.0 / F) as h:
        h.set_option(capi.OPT_MATVEC_PATH, path)
        h.set_data(X)
        for imax in (1, 2, 5, 10, 15, 17, 18, 19, 25, 60):
            a, it, delta = h.solver_cg(b, q, qa, imax, 0.0)
            ao, ito, do = port.solver_cg(X, b, q, qa, imax, 0.0, kernel, use_rows=True)
            print(path, imax, it, delta, "oracle", ito, do, float(np.abs(a).max()), flush=True)
