"""sddmm_reduce at the C3 weight-gradient shape for several replica counts (fixed vs per-slab cost)."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa
import torch_sputnik as ops
from oracle.csr import random_csr
dev = torch.device("cuda:0")
m = k = 4096
a = random_csr(m, k, 0.1, seed=0, values="pm1")
t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
for R, S in ((16, 512), (8, 512), (4, 512), (2, 512), (1, 512), (2, 256), (1, 2048), (1, 8192)):
    dy = torch.rand((R, m, S), device=dev); x = torch.rand((R, k, S), device=dev)
    fn = (lambda: ops.sddmm_reduce(m, k, ri, ro, ci, dy, x)) if R > 1 else (lambda: ops.sddmm(m, k, ri, ro, ci, dy[0], x[0]))
    for _ in range(3): fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(10): fn()
    e.record(); torch.cuda.synchronize()
    ms = s.elapsed_time(e) / 10
    print(f"R={R:2d} S={S:5d} K_eff={R*S:5d}  {ms:7.3f} ms  {2.0*a.nnz*R*S/ms/1e9:8.1f} TFLOP/s-ish(GF/ms)")
