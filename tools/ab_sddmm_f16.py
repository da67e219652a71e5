#!/usr/bin/env python
"""16-bit SDDMM at the C2 shape (4096 x 4096 mask, inner dimension 4096): the native sampled kernel on packed
operands against the widened fp32 path and the fp32 op itself, per density."""
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
    m = n = k = int(os.environ.get("TSP_S", "4096"))
    g = torch.Generator(device="cpu").manual_seed(1)
    lhs = torch.rand((m, k), generator=g).mul_(2).sub_(1).to(dev)
    rhs = torch.rand((n, k), generator=g).mul_(2).sub_(1).to(dev)
    for dens in (0.02, 0.05, 0.1, 0.15, 0.25, 0.35, 0.5):
        a = random_csr(m, n, dens, seed=0, values="pm1")
        t = lambda x: torch.from_numpy(x).to(dev)
        ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
        line = {"density": dens, "nnz": int(a.nnz)}
        line["f32"] = round(timeit(lambda: ops.sddmm(m, n, ri, ro, ci, lhs, rhs)), 4)
        for name, dt in (("f16", torch.float16), ("bf16", torch.bfloat16)):
            l, r = lhs.to(dt), rhs.to(dt)
            for label, knob in (("native", 2), ("widened", 1), ("auto", 0)):
                ops.set_variant("sddmm_b16", knob)
                line[f"{name} {label}"] = round(timeit(lambda: ops.sddmm(m, n, ri, ro, ci, l, r)), 4)
            ops.set_variant("sddmm_b16", 0)
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
