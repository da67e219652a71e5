"""sddmm_reduce / left_spmm at the C5 layer shapes (small masks, launch- and fixed-cost-bound): device time per
call from a back-to-back loop.  TSP_VARIANT forces an sddmm variant, TSP_SDDMM_BN the sampled tile width."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa
import torch_sputnik as ops
from oracle.csr import random_csr
dev = torch.device("cuda:0")
if os.environ.get("TSP_VARIANT"):
    ops.set_variant("sddmm", int(os.environ["TSP_VARIANT"]))
if os.environ.get("TSP_SDDMM_BN"):
    ops.set_variant("sddmm_bn", int(os.environ["TSP_SDDMM_BN"]))
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)


def timeit(fn, reps=50):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps): fn()
    e.record(); torch.cuda.synchronize()
    return s.elapsed_time(e) / reps * 1e3


for (m, k) in ((512, 512), (2048, 512), (512, 2048)):
    a = random_csr(m, k, 0.1, seed=0, values="pm1")
    v, ri, ro, ci = t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices)
    R, S = 4, 512
    dy = torch.rand((R, m, S), device=dev); x = torch.rand((R, k, S), device=dev)
    us_w = timeit(lambda: ops.sddmm_reduce(m, k, ri, ro, ci, dy, x))
    us_f = timeit(lambda: ops.left_spmm(m, k, v, ri, ro, ci, x))
    print(f"weight {m}x{k} nnz={a.nnz}: wgrad sddmm_reduce {us_w:7.1f} us   fwd left_spmm {us_f:7.1f} us")
