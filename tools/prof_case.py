"""Run one op of the C2 workload a few times (short command line for ncu captures).
usage: python tools/prof_case.py spmm|sddmm|softmax|transpose DENSITY [REPS] [N]"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from oracle.csr import dense_operand, random_csr  # noqa: E402

op, d = sys.argv[1], float(sys.argv[2])
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
n = int(sys.argv[4]) if len(sys.argv) > 4 else 4096
dim = 4096
if os.environ.get("TSP_VARIANT"):
    ops.set_variant(op if op in ("spmm", "sddmm") else "spmm", int(os.environ["TSP_VARIANT"]))
for knob in ("sddmm_bn", "sddmm_p", "spmm_scb", "sddmm_tiled3"):
    if os.environ.get("TSP_" + knob.upper()):
        ops.set_variant(knob, int(os.environ["TSP_" + knob.upper()]))
dev = torch.device("cuda:0")
a = random_csr(dim, dim, d, seed=0, values="pm1")
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
v, ri, ro, ci = t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices)
b = t(dense_operand((dim, n), 1, "pm1"))
lhs, rhs = t(dense_operand((dim, n), 2, "pm1")), t(dense_operand((dim, n), 3, "pm1"))
if os.environ.get("TSP_DTYPE") in ("f16", "bf16"):   # 16-bit storage, fp32 arithmetic
    dt = torch.float16 if os.environ["TSP_DTYPE"] == "f16" else torch.bfloat16
    v, b, lhs, rhs = v.to(dt), b.to(dt), lhs.to(dt), rhs.to(dt)


def call():
    if op == "spmm":
        ops.spmm(dim, dim, v, ri, ro, ci, b)
    elif op == "sddmm":
        ops.sddmm(dim, dim, ri, ro, ci, lhs, rhs)
    elif op == "softmax":
        ops.sparse_softmax(v, ri, ro, ci)
    else:
        ops.csr_transpose(dim, dim, v, ro, ci)


call()                       # plans the topology; its verdict is on the host after the synchronize
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStart()      # `ncu --profile-from-start off`: steady-state launches only
ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
ev[0].record()
for i in range(reps):
    call()
    ev[i + 1].record()
torch.cuda.synchronize()
torch.cuda.cudart().cudaProfilerStop()
print(op, d, n, [round(ev[i].elapsed_time(ev[i + 1]), 4) for i in range(reps)], "ms")
