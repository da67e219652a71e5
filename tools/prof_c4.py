"""One op at the attention shape (s=4096, d=64, 90 % sparse mask), few replicas -- for ncu captures.
usage: python tools/prof_c4.py spmm|sddmm|softmax|attn [REPLICAS] [REPS]   (TSP_VARIANT forces a kernel variant, TSP_DTYPE=f16)"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from oracle.csr import random_csr  # noqa: E402

op = sys.argv[1]
R = int(sys.argv[2]) if len(sys.argv) > 2 else 16
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
if os.environ.get("TSP_VARIANT") and op != "attn":
    ops.set_variant(op, int(os.environ["TSP_VARIANT"]))
for knob in ("spmm_warps", "sddmm_warps", "attn_warps", "band_warps"):
    if os.environ.get("TSP_" + knob.upper()):
        ops.set_variant(knob, int(os.environ["TSP_" + knob.upper()]))
dev = torch.device("cuda:0")
s, d = 4096, 64
a = random_csr(s, s, 0.1, seed=0, values="pm1", round_to=4)
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
q, k, v = (torch.rand((R, s, d), device=dev) * 2 - 1 for _ in range(3))
p = torch.rand((R, a.nnz), device=dev)
if os.environ.get("TSP_DTYPE") == "f16":
    q, k, v, p = q.half(), k.half(), v.half(), p.half()


def call():
    if op == "spmm":
        ops.spmm(s, s, p, ri, ro, ci, v)
    elif op == "attn":
        ops.sparse_attention(s, s, ri, ro, ci, q, k, v, 0.125)
    elif op == "softmax":
        ops.sparse_softmax_scaled(p, ri, ro, ci, 0.125)
    else:
        ops.sddmm(s, s, ri, ro, ci, q, k)


call()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()      # `ncu --profile-from-start off`: steady-state launches only
ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
ev[0].record()
for i in range(reps):
    call()
    ev[i + 1].record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(op, "R=%d" % R, [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(reps)], "ms")
