#!/usr/bin/env python
"""A/B at the C4 attention shape (R replicas x 4096^2 mask at 90 % sparsity x d = 64): the L1-gather
kernels against the shared-memory band kernels, fp32 and 16-bit storage.  One line per variant."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from oracle.csr import random_csr  # noqa: E402  (input generator only)


def timeit(fn, reps=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def main():
    dev = torch.device("cuda:0")
    r = int(os.environ.get("TSP_R", "128"))
    s, d = int(os.environ.get("TSP_S", "4096")), int(os.environ.get("TSP_D", "64"))
    dens = float(os.environ.get("TSP_DENSITY", "0.1"))
    a = random_csr(s, s, dens, seed=0, values="pm1", round_to=4)
    t = lambda x: torch.from_numpy(x).to(dev)
    ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
    g = torch.Generator(device="cpu").manual_seed(1)
    q, k, v = (torch.rand((r, s, d), generator=g).mul_(2).sub_(1).to(dev) for _ in range(3))
    for name, dt in (("f32", torch.float32), ("f16", torch.float16)):
        qd, kd, vd = q.to(dt), k.to(dt), v.to(dt)
        sc = ops.sddmm(s, s, ri, ro, ci, qd, kd)
        pr = ops.sparse_softmax_scaled(sc, ri, ro, ci, 0.125)
        res = {}
        for label, var, bw in (("rows", 5, 0), ("band16", 6, 16), ("band32", 6, 32)):
            ops.set_variant("sddmm", var)
            ops.set_variant("band_warps", bw)
            res["sddmm " + label] = round(timeit(lambda: ops.sddmm(s, s, ri, ro, ci, qd, kd)), 4)
        for label, var, bw in (("gather", 1, 0), ("rowsplit", 5, 0), ("band16", 6, 16), ("band32", 6, 32)):
            if name != "f32" and label == "gather":
                continue
            ops.set_variant("spmm", var)
            ops.set_variant("band_warps", bw)
            res["spmm " + label] = round(timeit(lambda: ops.spmm(s, s, pr, ri, ro, ci, vd)), 4)
        ops.set_variant("sddmm", 0)
        ops.set_variant("spmm", 0)
        ops.set_variant("band_warps", 0)
        res["softmax"] = round(timeit(lambda: ops.sparse_softmax_scaled(sc, ri, ro, ci, 0.125)), 4)
        res["sddmm auto"] = round(timeit(lambda: ops.sddmm(s, s, ri, ro, ci, qd, kd)), 4)
        res["spmm auto"] = round(timeit(lambda: ops.spmm(s, s, pr, ri, ro, ci, vd)), 4)
        for w in (32, 16, 8):
            ops.set_variant("attn_warps", w)
            res[f"fused attention w{w}"] = round(timeit(lambda: ops.sparse_attention(s, s, ri, ro, ci, qd, kd, vd, 0.125)), 4)
        ops.set_variant("attn_warps", 0)
        print(json.dumps({"dtype": name, "replicas": r, "seq": s, "d": d, "density": dens, "nnz": int(a.nnz), "ms": res}))


if __name__ == "__main__":
    main()
