import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from plssvm_sparse_fp22_b200 import capi, datagen
n, F = 262144, 20000
rp, ci, v, y = datagen.make_sparse_csr(n, F, 1e-2, seed=45, dtype=np.float32)
v = datagen.fp22_round(v)
prp, packed = datagen.pack_fp22_csr(rp, ci, v)
b = (y[:-1] - y[-1]).astype(np.float32)
print("free before", torch.cuda.mem_get_info())
h = capi.Handle(np.float32, "polynomial", 3, 1.0 / F, 0.0, 1.0)
h.set_option(capi.OPT_SPARSE_TILES, int(sys.argv[1]) if len(sys.argv) > 1 else 0)
h.set_data_csr(prp, None, packed, F, value_format=capi.VAL_FP22)
print("free after set_data", torch.cuda.mem_get_info())
h.generate_q(); QA = h.kxx_last() + 1.0
t0 = time.time()
try:
    a, it, d = h.solver_cg(b, None, QA, 3, 0.0)
    torch.cuda.synchronize()
    st = h.stats()
    print("solve ok", time.time() - t0, "dense bytes", st.dense_cache_bytes, "entries", st.sparse_cache_entries, "build ms", st.cache_build_ms, "mv ms", st.matvec_ms, st.matvecs)
except Exception as e:
    print("ERR", e)
print("free after solve", torch.cuda.mem_get_info())
print("atoms gops", h.probe_smem_atomic_gops(), "fma f32", h.probe_fma_tflops(np.float32), "fma f64", h.probe_fma_tflops(np.float64))
