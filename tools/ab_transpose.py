#!/usr/bin/env python
"""A/B of the two csr_transpose paths through the raw C-ABI (no plan cache in the way): the counting
path (count -> offsets -> place, 3 launches) against the multi-pass radix path, bit-compared with each
other.  One JSON line per shape."""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
from torch_sputnik_b200 import cabi  # noqa: E402
from oracle.csr import random_csr  # noqa: E402  (input generator only)

SHAPES = [("c3_weight 3072x768 d=0.1", 3072, 768, 0.1), ("c3_weight_t 768x3072 d=0.1", 768, 3072, 0.1),
          ("c4_mask 4096x4096 d=0.1", 4096, 4096, 0.1), ("c2 8192x8192 d=0.05", 8192, 8192, 0.05),
          ("c2 8192x8192 d=0.5", 8192, 8192, 0.5), ("wide 512x40000 d=0.02", 512, 40000, 0.02)]


def main():
    dev = torch.device("cuda:0")
    lib = cabi.load()
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for name, m, n, dens in SHAPES:
        a = random_csr(m, n, dens, seed=3, values="pm1")
        v = torch.from_numpy(a.values).to(dev)
        ro = torch.from_numpy(a.row_offsets).to(dev)
        ci = torch.from_numpy(a.column_indices).to(dev)
        line = {"shape": name, "nnz": int(a.nnz)}
        outs = {}
        for label, variant in (("counting", 0), ("radix", 1)):
            lib.tsp_set_variant(b"transpose", variant)
            size = lib.tsp_csr_transpose_workspace(m, n, a.nnz)
            big = size
            ws = torch.empty(big, dtype=torch.uint8, device=dev)
            vt = torch.empty_like(v)
            rot = torch.empty(n + 1, dtype=torch.int32, device=dev)
            cit = torch.empty(a.nnz, dtype=torch.int32, device=dev)
            perm = torch.empty(a.nnz, dtype=torch.int32, device=dev)

            def run():
                rc = lib.tsp_csr_transpose_f32(m, n, a.nnz, 1, p(v), p(ro), p(ci), p(vt), p(rot), p(cit), p(perm),
                                               p(ws), big, stream)
                assert rc == 0, cabi.last_error()

            for _ in range(3):
                run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20):
                run()
            e1.record()
            torch.cuda.synchronize()
            line[label + "_us"] = round(e0.elapsed_time(e1) / 20 * 1e3, 1)
            outs[label] = (vt.clone(), rot.clone(), cit.clone(), perm.clone())
        lib.tsp_set_variant(b"transpose", 0)
        line["identical"] = all(torch.equal(x, y) for x, y in zip(outs["counting"], outs["radix"]))
        print(json.dumps(line), flush=True)


if __name__ == "__main__":
    main()
