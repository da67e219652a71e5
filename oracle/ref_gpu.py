"""ctypes front end of oracle/_ref/libsputnik_ref.so -- the reference's OWN kernels (Sputnik,
compiled unmodified for sm_100a) and cuSPARSE csr2csc, on torch CUDA tensors.

TEST / MEASUREMENT INFRASTRUCTURE ONLY.  Used (a) to generate tests/golden/*.npz on a B200
(tests/golden/make_golden_gpu.py), (b) as a live GPU-side parity reference in ``-m gpu`` tests,
(c) as "the reference's kernel on the same box" timing in bench.py.  Never on the product path.
"""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "_ref", "libsputnik_ref.so")
_lib = None


def available() -> bool:
    return os.path.exists(LIB_PATH)


def _load():
    global _lib
    if _lib is None:
        lib = ctypes.CDLL(LIB_PATH)
        i, vp, sz = ctypes.c_int, ctypes.c_void_p, ctypes.c_size_t
        lib.ref_spmm.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.ref_spmm_half.argtypes = [i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.ref_sddmm.argtypes = [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
        lib.ref_sparse_softmax.argtypes = [i, i, i, vp, vp, vp, vp, vp, vp]
        lib.ref_csr_transpose.argtypes = [i, i, i, vp, vp, vp, vp, vp, vp, vp]
        lib.ref_csr_transpose_workspace.argtypes = [i, i, i, vp, vp, vp, vp, vp, vp]
        lib.ref_csr_transpose_workspace.restype = sz
        lib.ref_csr_transpose_ws.argtypes = [i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        _lib = lib
    return _lib


def _stream():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else ctypes.c_void_p(0)


def spmm(m, k, values, row_indices, row_offsets, column_indices, dense, bias=None,
         values_shared=False):
    """src/spmm_cuda.cu:9-60 (values_shared=False) / src/left_replicated_spmm.cu:8-44 (True)."""
    import torch
    lib = _load()
    reps = dense.shape[0] if dense.dim() == 3 else 1
    n = dense.shape[-1]
    nnz = column_indices.shape[0]
    out = torch.zeros((reps, m, n), dtype=torch.float32, device=dense.device)
    rc = lib.ref_spmm(m, k, n, nnz, reps, int(values_shared), _p(row_indices), _p(values),
                      _p(row_offsets), _p(column_indices), _p(dense), _p(bias), _p(out), _stream())
    if rc != 0:
        raise RuntimeError(f"ref_spmm failed: cudaError {rc}")
    return out if dense.dim() == 3 else out[0]


def spmm_half(m, k, values, row_indices, row_offsets, column_indices, dense, bias=None):
    """sputnik::CudaSpmm / CudaSpmmBiasRelu, half2 / short2 overloads (cuda_spmm.cu.cc:511-570).
    values, dense: float16; column_indices: int32 (narrowed to the int16 the kernel reads).  The
    contract of cuda_spmm.h:75-113 is asserted: even nonzero count in every row, n % 4 == 0 (half4 dense loads, cuda_spmm.cu.cc:548-553), k < 32768."""
    import torch
    lib = _load()
    n = dense.shape[-1]
    nnz = column_indices.shape[0]
    assert values.dtype == torch.float16 and dense.dtype == torch.float16 and dense.dim() == 2
    assert n % 4 == 0 and k < 32768 and bool((row_offsets % 2 == 0).all()), "cuda_spmm.h:75-113"
    cols16 = column_indices.to(torch.int16).contiguous()
    out = torch.zeros((m, n), dtype=torch.float16, device=dense.device)
    rc = lib.ref_spmm_half(m, k, n, nnz, _p(row_indices), _p(values), _p(row_offsets), _p(cols16), _p(dense),
                           _p(bias), _p(out), _stream())
    if rc != 0:
        raise RuntimeError(f"ref_spmm_half failed: cudaError {rc}")
    return out


def sddmm(m, n, row_indices, row_offsets, column_indices, lhs, rhs):
    import torch
    lib = _load()
    reps = lhs.shape[0] if lhs.dim() == 3 else 1
    k = lhs.shape[-1]
    nnz = column_indices.shape[0]
    out = torch.zeros((reps, nnz), dtype=torch.float32, device=lhs.device)
    rc = lib.ref_sddmm(m, k, n, nnz, reps, _p(row_indices), _p(row_offsets), _p(column_indices),
                       _p(lhs), _p(rhs), _p(out), _stream())
    if rc != 0:
        raise RuntimeError(f"ref_sddmm failed: cudaError {rc}")
    return out if lhs.dim() == 3 else out[0]


def sparse_softmax(values, row_indices, row_offsets, column_indices):
    import torch
    lib = _load()
    reps = values.shape[0] if values.dim() == 2 else 1
    nnz = column_indices.shape[0]
    out = torch.zeros_like(values)
    rc = lib.ref_sparse_softmax(row_indices.shape[0], nnz, reps, _p(values), _p(row_indices),
                                _p(row_offsets), _p(column_indices), _p(out), _stream())
    if rc != 0:
        raise RuntimeError(f"ref_sparse_softmax failed: cudaError {rc}")
    return out


def csr_transpose(m, n, values, row_offsets, column_indices):
    """cuSPARSE csr2cscEx2 ALG1 exactly as src/transpose_cuda.cu:90-99 calls it."""
    import torch
    lib = _load()
    nnz = column_indices.shape[0]
    vt = torch.zeros(nnz, dtype=torch.float32, device=values.device)
    rot = torch.zeros(n + 1, dtype=torch.int32, device=values.device)
    cit = torch.zeros(nnz, dtype=torch.int32, device=values.device)
    rc = lib.ref_csr_transpose(m, n, nnz, _p(values), _p(row_offsets), _p(column_indices), _p(vt),
                               _p(rot), _p(cit), _stream())
    if rc != 0:
        raise RuntimeError(f"ref_csr_transpose failed: {rc}")
    return vt, rot, cit


class TransposeTimer:
    """cuSPARSE csr2csc with a preallocated workspace (for timing without cudaMalloc)."""

    def __init__(self, m, n, values, row_offsets, column_indices):
        import torch
        self.lib = _load()
        self.args = (m, n, column_indices.shape[0])
        nnz = self.args[2]
        self.inputs = (values, row_offsets, column_indices)
        self.vt = torch.zeros(nnz, dtype=torch.float32, device=values.device)
        self.rot = torch.zeros(n + 1, dtype=torch.int32, device=values.device)
        self.cit = torch.zeros(nnz, dtype=torch.int32, device=values.device)
        size = self.lib.ref_csr_transpose_workspace(m, n, nnz, _p(values), _p(row_offsets),
                                                    _p(column_indices), _p(self.vt), _p(self.rot),
                                                    _p(self.cit))
        self.ws = torch.zeros(max(int(size), 4), dtype=torch.uint8, device=values.device)

    def __call__(self):
        m, n, nnz = self.args
        v, ro, ci = self.inputs
        rc = self.lib.ref_csr_transpose_ws(m, n, nnz, _p(v), _p(ro), _p(ci), _p(self.vt),
                                           _p(self.rot), _p(self.cit), _p(self.ws), _stream())
        if rc != 0:
            raise RuntimeError(f"ref_csr_transpose_ws failed: {rc}")
