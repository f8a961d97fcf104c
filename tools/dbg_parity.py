"""Diagnostic (multi-GPU): alpha of K fixed CG iterations on N ranks vs one rank, K = 1..20 (configs[1] data)."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch, torch.distributed as dist
from plssvm_sparse_fp22_b200 import capi, datagen

rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1")); local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
n, F = int(sys.argv[1]) if len(sys.argv) > 1 else 65536, int(sys.argv[2]) if len(sys.argv) > 2 else 1024
X, y = datagen.make_dense(n, F, seed=43, problem="blobs", dtype=np.float32)
b = (y[:-1] - y[-1]).astype(np.float32)
def uid():
    idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idbuf.copy_(torch.frombuffer(bytearray(capi.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(idbuf, 0)
    return bytes(idbuf.cpu().numpy().tobytes())
hN = capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0, device=local)
if world > 1:
    hN.comm_init(uid(), rank, world)
hN.set_data(X); hN.generate_q(); QA = hN.kxx_last() + 1.0
h1 = None
if rank == 0:
    h1 = capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0, device=local)
    h1.set_data(X); h1.generate_q()
    hs = capi.Handle(np.float32, "rbf", 3, 1.0 / F, 0.0, 1.0, device=local)
    hs.set_option(capi.OPT_MATVEC_PATH, capi.PATH_SIMT)
    hs.set_data(X[:8192] if n > 8192 else X)
for K in (1, 2, 3, 5, 10, 20):
    aN, itN, dN = hN.solver_cg(b, None, QA, K, 0.0)
    if rank == 0:
        a1, it1, d1 = h1.solver_cg(b, None, QA, K, 0.0)
        e = np.linalg.norm(aN.astype(np.float64) - a1) / np.linalg.norm(a1)
        print(f"K={K:3d} delta_N={dN:.4e} delta_1={d1:.4e} |alpha_1|={np.linalg.norm(a1):.4e} relerr={e:.3e}", flush=True)
    if world > 1:
        dist.barrier()
hN.close()
if world > 1:
    dist.destroy_process_group()
