"""Generate tests/golden/ref_b200.npz ON A B200: outputs of the reference's OWN kernels.

Run (from the repo root, on the GPU box; oracle/_ref/libsputnik_ref.so is built here in the
container by `make -C oracle ref` from the unmodified sources under /root/reference and travels
with the snapshot):

    python tests/golden/make_golden_gpu.py gpurun_out/ref_b200.npz

then copy gpurun_out/ref_b200.npz to tests/golden/ref_b200.npz and commit it.  Every case stores
its inputs (so nothing depends on RNG stability) and the reference's outputs:

  spmm_*      sputnik::CudaSpmm / CudaSpmmBiasRelu   (src/spmm_cuda.cu:48-57 host loop)
  sddmm_*     sputnik::CudaSddmm                     (src/sddmm_cuda.cu:45-54)
  softmax_*   sputnik::SparseSoftmax                 (src/softmax_cuda.cu:35-43)
  transpose_* cusparseCsr2cscEx2(ALG1)               (src/transpose_cuda.cu:90-99)

`python tests/golden/make_golden_gpu.py --half gpurun_out/ref_b200_f16.npz` writes the second file,
tests/golden/ref_b200_f16.npz:

  spmmh_*     sputnik::CudaSpmm / CudaSpmmBiasRelu, half2 / short2 overloads (cuda_spmm.cu.cc:511-570):
              fp16 values and dense operand, fp32 accumulation, fp16 result; rows padded to an even
              nonzero count as spmm_test.cu.cc:172-176 does (here: the odd entry is dropped)
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import ref_gpu  # noqa: E402
from oracle.csr import dense_operand, random_csr  # noqa: E402

SPMM = [  # name, m, k, n, nnz/density, values, bias
    ("t72", 72, 64, 72, dict(density=0.1, round_to=4), "u01", False),       # tests/test_spmm.py:13
    ("b64", 64, 64, 64, dict(density=0.25), "u01", False),                  # README.md:35 / BASELINE configs[0]
    ("p0", 4, 32, 32, dict(nnz=32), "pm1", False),                          # spmm_test.cu.cc:80-95
    ("p1", 4, 32, 48, dict(nnz=128), "pm1", False),
    ("p2", 7, 96, 40, dict(nnz=333), "pm1", False),
    ("p3", 11, 55, 8, dict(nnz=293), "pm1", False),
    ("p4", 4, 128, 128, dict(nnz=512), "pm1", False),
    ("p5", 64, 512, 512, dict(nnz=16384), "pm1", False),
    ("bias", 72, 64, 72, dict(density=0.3), "pm1", True),                   # cuda_spmm.cu.cc:267-277
    ("odd", 37, 53, 33, dict(density=0.3), "pm1", False),
]
SDDMM = [  # name, m, k, n, nnz/density      (sddmm_test.cu.cc:76-91)
    ("t72", 72, 64, 72, dict(density=1.0)),                                 # tests/test_sddmm.py:16
    ("p0", 1, 32, 32, dict(nnz=32)), ("p1", 1, 64, 32, dict(nnz=32)), ("p2", 1, 32, 64, dict(nnz=64)),
    ("p3", 1, 32, 64, dict(nnz=32)), ("p4", 1, 32, 64, dict(nnz=16)), ("p5", 2, 32, 32, dict(nnz=64)),
    ("p7", 1, 28, 32, dict(nnz=32)), ("p6s", 256, 80, 256, dict(nnz=8192)),   # scaled-down 1k rnn case
    ("att", 256, 64, 256, dict(density=0.1)),
]
SOFTMAX = [("t72", 72, 72, 0.1), ("long", 16, 4096, 0.5), ("wide", 128, 4096, 0.1)]
TRANSPOSE = [("kat4", 4, 4, 0.5), ("t72", 72, 64, 0.1), ("r300", 300, 1000, 0.02), ("tall", 1000, 37, 0.3),
             ("full", 64, 64, 1.0), ("wide", 5, 70000, 0.001), ("big", 512, 512, 0.1)]


SPMM_HALF = [  # name, m, k, n, nnz/density, bias      (shapes: spmm_test.cu.cc:80-95; n even, k < 32768)
    ("p0", 4, 32, 32, dict(nnz=32), False),
    ("p1", 4, 32, 48, dict(nnz=128), False),
    ("p2", 7, 96, 40, dict(nnz=333), False),
    ("p3", 11, 55, 8, dict(nnz=293), False),
    ("p4", 4, 128, 128, dict(nnz=512), False),
    ("p5", 64, 512, 512, dict(nnz=16384), False),
    ("b1k", 1024, 1024, 128, dict(nnz=131072), False),
    ("att", 256, 256, 64, dict(density=0.1), False),
    ("bias", 72, 64, 72, dict(density=0.3), True),
    ("n36", 37, 53, 36, dict(density=0.3), False),   # n % 4 == 0: the half4 dense loads of the fall-through config
]


def main_half(out_path):
    from oracle.csr import even_rows
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    g = {}
    for name, m, k, n, kw, with_bias in SPMM_HALF:
        a = even_rows(random_csr(m, k, seed=500 + len(g), values="pm1", **kw))
        vals = a.values.astype(np.float16)
        b = dense_operand((k, n), 600 + len(g), "pm1").astype(np.float16)
        bias = dense_operand((m,), 700 + len(g), "pm1") if with_bias else None
        out = ref_gpu.spmm_half(m, k, t(vals), t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(b),
                                bias=None if bias is None else t(bias))
        p = f"spmmh_{name}_"
        g[p + "shape"] = np.array([m, k, n], dtype=np.int32)
        g[p + "values"], g[p + "row_indices"] = vals, a.row_indices
        g[p + "row_offsets"], g[p + "column_indices"] = a.row_offsets, a.column_indices
        g[p + "dense"], g[p + "out"] = b, out.cpu().numpy()
        if bias is not None:
            g[p + "bias"] = bias
    torch.cuda.synchronize()
    g["meta"] = np.array([torch.cuda.get_device_name(0), torch.__version__, torch.version.cuda or ""])
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **g)
    print("wrote", out_path, os.path.getsize(out_path), "bytes;", len(g), "arrays")


def main(out_path):
    dev = torch.device("cuda:0")
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    g = {}
    for name, m, k, n, kw, dist, with_bias in SPMM:
        a = random_csr(m, k, seed=len(g), values=dist, **kw)
        b = dense_operand((k, n), 100 + len(g), dist)
        bias = dense_operand((m,), 200 + len(g), "pm1") if with_bias else None
        out = ref_gpu.spmm(m, k, t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(b),
                           bias=None if bias is None else t(bias))
        p = f"spmm_{name}_"
        g[p + "shape"] = np.array([m, k, n], dtype=np.int32)
        g[p + "values"], g[p + "row_indices"] = a.values, a.row_indices
        g[p + "row_offsets"], g[p + "column_indices"] = a.row_offsets, a.column_indices
        g[p + "dense"], g[p + "out"] = b, out.cpu().numpy()
        if bias is not None:
            g[p + "bias"] = bias
    for name, m, k, n, kw in SDDMM:
        a = random_csr(m, n, seed=len(g), values="pm1", **kw)
        lhs, rhs = dense_operand((m, k), 300 + len(g), "pm1"), dense_operand((n, k), 400 + len(g), "pm1")
        out = ref_gpu.sddmm(m, n, t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(lhs), t(rhs))
        p = f"sddmm_{name}_"
        g[p + "shape"] = np.array([m, k, n], dtype=np.int32)
        g[p + "row_indices"], g[p + "row_offsets"], g[p + "column_indices"] = a.row_indices, a.row_offsets, a.column_indices
        g[p + "lhs"], g[p + "rhs"], g[p + "out"] = lhs, rhs, out.cpu().numpy()
    for name, m, n, d in SOFTMAX:
        a = random_csr(m, n, d, seed=len(g), values="pm1")
        x = (a.values * 8).astype(np.float32)
        out = ref_gpu.sparse_softmax(t(x), t(a.row_indices), t(a.row_offsets), t(a.column_indices))
        p = f"softmax_{name}_"
        g[p + "values"], g[p + "row_indices"], g[p + "row_offsets"] = x, a.row_indices, a.row_offsets
        g[p + "column_indices"], g[p + "out"] = a.column_indices, out.cpu().numpy()
    for name, m, n, d in TRANSPOSE:
        a = random_csr(m, n, d, seed=len(g), values="pm1")
        vt, rot, cit = ref_gpu.csr_transpose(m, n, t(a.values), t(a.row_offsets), t(a.column_indices))
        p = f"transpose_{name}_"
        g[p + "shape"] = np.array([m, n], dtype=np.int32)
        g[p + "values"], g[p + "row_offsets"], g[p + "column_indices"] = a.values, a.row_offsets, a.column_indices
        g[p + "values_t"], g[p + "row_offsets_t"] = vt.cpu().numpy(), rot.cpu().numpy()
        g[p + "column_indices_t"] = cit.cpu().numpy()
    torch.cuda.synchronize()
    g["meta"] = np.array([torch.cuda.get_device_name(0), torch.__version__, torch.version.cuda or ""])
    os.makedirs(os.path.dirname(os.path.abspath(out_path)), exist_ok=True)
    np.savez_compressed(out_path, **g)
    print("wrote", out_path, os.path.getsize(out_path), "bytes;", len(g), "arrays")


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--half":
        main_half(sys.argv[2] if len(sys.argv) > 2 else os.path.join(ROOT, "gpurun_out", "ref_b200_f16.npz"))
    else:
        main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "ref_b200.npz"))
