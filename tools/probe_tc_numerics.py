"""GPU probe (design input, not product): how accurate is a Gram matrix G = X X^T built from tensor-core products
with fp32 accumulation on B200, against an fp64 reference?  Variants:
  fp32      : torch.mm fp32 with TF32 disabled (SIMT FFMA)
  tf32      : torch.mm fp32 with TF32 enabled (one tensor pass)
  f16x2_3p  : x = hi + lo (two fp16), hi*hi + hi*lo + lo*hi — three fp16 tensor passes summed in fp32
  f16x2_cat : same three products as ONE GEMM with K concatenated ([hi|hi|lo] x [hi|lo|hi]) — single accumulator
  bf16x3_6p : three bf16 terms, six products
Reports max / rms error of G normalised by sqrt(G_ii G_jj) and the mean signed error (accumulator truncation bias),
then the induced relative error of an RBF kernel entry exp(-gamma d2), gamma = 1/F.
"""
import json
import sys

import numpy as np
import torch

sys.path.insert(0, ".")
from plssvm_sparse_fp22_b200 import datagen  # noqa: E402


def mm32(a, b):
    try:
        return torch.mm(a, b.t(), out_dtype=torch.float32)
    except TypeError:
        return torch.mm(a, b.t()).float()


def main():
    n, F = 4096, 1024
    X, _ = datagen.make_dense(n, F, seed=43, problem="blobs", dtype=np.float32)
    x32 = torch.from_numpy(X).cuda()
    x64 = x32.double()
    G64 = x64 @ x64.t()
    nrm = torch.sqrt(torch.diag(G64))
    scale = nrm[:, None] * nrm[None, :]
    gamma = 1.0 / F
    res = {}

    def report(name, G):
        e = (G.double() - G64) / scale
        d2err = 2.0 * (G.double() - G64).abs() * gamma  # |delta arg| of exp(-gamma (ni+nj-2G))
        res[name] = {"max_rel": float(e.abs().max()), "rms_rel": float(e.pow(2).mean().sqrt()), "mean_signed": float(e.mean()),
                     "rbf_rel_err_max": float(d2err.max()), "rbf_rel_err_rms": float(d2err.pow(2).mean().sqrt())}

    torch.backends.cuda.matmul.allow_tf32 = False
    report("fp32", x32 @ x32.t())
    torch.backends.cuda.matmul.allow_tf32 = True
    report("tf32", x32 @ x32.t())
    torch.backends.cuda.matmul.allow_tf32 = False

    amax = float(x32.abs().max())
    s = 2.0 ** (14 - int(np.ceil(np.log2(amax))))  # place max |x| in [2^13, 2^14): fp16 keeps 11 bits down to 2^-14
    xs = x32 * s
    hi = xs.half()
    lo = (xs - hi.float()).half()
    inv = 1.0 / (s * s)
    report("f16x2_3p", (mm32(hi, hi) + (mm32(hi, lo) + mm32(lo, hi))) * inv)
    A = torch.cat([hi, hi, lo], dim=1).contiguous()
    B = torch.cat([hi, lo, hi], dim=1).contiguous()
    report("f16x2_cat", mm32(A, B) * inv)
    report("f16_1p", mm32(hi, hi) * inv)

    b0 = x32.bfloat16()
    r1 = x32 - b0.float()
    b1 = r1.bfloat16()
    b2 = (r1 - b1.float()).bfloat16()
    G = mm32(b0, b0) + (mm32(b0, b1) + mm32(b1, b0)) + (mm32(b1, b1) + mm32(b0, b2) + mm32(b2, b0))
    report("bf16x3_6p", G)
    report("bf16x2_3p", mm32(b0, b0) + (mm32(b0, b1) + mm32(b1, b0)))
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
