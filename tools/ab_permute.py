#!/usr/bin/env python
"""permute_values at the C4 shape ([R, nnz] probabilities through the transpose permutation of a 4096^2
mask at 90 % sparsity) and the C3 weight: run length per CTA (items x threads) against the element-per-
thread launch.  GB/s counts 8 B per value + 4 B per index."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from oracle.csr import random_csr  # noqa: E402  (input generator only)

CFGS = [108, 816, 832]


def timeit(fn, reps=20, warm=3):
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
    for name, m, n, dens, r, dt in (("c4 f32", 4096, 4096, 0.1, 128, torch.float32),
                                    ("c4 f16", 4096, 4096, 0.1, 128, torch.float16),
                                    ("c4 d=0.3 f32", 4096, 4096, 0.3, 32, torch.float32),
                                    ("c3 weight", 3072, 768, 0.1, 1, torch.float32)):
        a = random_csr(m, n, dens, seed=0, values="pm1")
        ro = torch.from_numpy(a.row_offsets).to(dev)
        ci = torch.from_numpy(a.column_indices).to(dev)
        rot, cit, perm, rit = ops.transposed_topology(m, n, ro, ci)
        v = torch.rand((r, a.nnz), device=dev).to(dt)
        ref = v[:, perm.long()]
        line = {"case": name, "nnz": int(a.nnz), "replicas": r}
        bytes_moved = r * a.nnz * 2 * v.element_size() + 4 * a.nnz
        ops.set_variant("permute_tiled", 1)
        for cfg in CFGS:
            ops.set_variant("permute", cfg)
            out = ops.permute_values(v, perm)
            assert torch.equal(out, ref)
            ms = timeit(lambda: ops.permute_values(v, perm))
            line["%dx%d" % (cfg // 100, (cfg % 100) * 32)] = [round(ms * 1e3, 1), round(bytes_moved / ms / 1e6)]
        ops.set_variant("permute", 0)
        ops.set_variant("permute_tiled", 1)
        for label, knob in (("gather", 1), ("tiled", 2), ("default", 0)):
            ops.set_variant("permute_tiled", knob)
            out = ops.permute_values(v, perm)
            assert torch.equal(out, ref), label
            ms = timeit(lambda: ops.permute_values(v, perm))
            line[label] = [round(ms * 1e3, 1), round(bytes_moved / ms / 1e6)]
        ops.set_variant("permute_tiled", 0)
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
