"""ctypes front end of oracle.c plus the dense torch references (TEST INFRASTRUCTURE)."""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liboracle.so")
_lib = None

_f32p = np.ctypeslib.ndpointer(dtype=np.float32, flags="C_CONTIGUOUS")
_i32p = np.ctypeslib.ndpointer(dtype=np.int32, flags="C_CONTIGUOUS")


def build(force: bool = False) -> str:
    """Compile oracle.c -> liboracle.so (gcc).  Building the checker is not using it."""
    src = os.path.join(_HERE, "oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "liboracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


def _load():
    global _lib
    if _lib is None:
        build()
        lib = ctypes.CDLL(_LIB_PATH)
        c_int, c_float, vp = ctypes.c_int, ctypes.c_float, ctypes.c_void_p
        lib.oracle_spmm.argtypes = [c_int, c_int, c_int, _f32p, _i32p, _i32p, _f32p, vp, c_int, _f32p]
        lib.oracle_spmm_abs.argtypes = [c_int, c_int, c_int, _f32p, _i32p, _i32p, _f32p, _f32p]
        lib.oracle_sddmm.argtypes = [c_int, c_int, c_int, _i32p, _i32p, _f32p, _f32p, _f32p]
        lib.oracle_sddmm_abs.argtypes = [c_int, c_int, c_int, _i32p, _i32p, _f32p, _f32p, _f32p]
        lib.oracle_csr_softmax.argtypes = [c_int, _f32p, _i32p, c_float, _f32p]
        lib.oracle_csr_softmax_f32.argtypes = [c_int, _f32p, _i32p, _f32p]
        lib.oracle_csr_softmax_backward.argtypes = [c_int, _f32p, _f32p, _i32p, c_float, _f32p]
        lib.oracle_csr_transpose.argtypes = [c_int, c_int, c_int, _f32p, _i32p, _i32p, _f32p, _i32p,
                                             _i32p, _i32p]
        for name in ("oracle_spmm", "oracle_spmm_abs", "oracle_sddmm", "oracle_sddmm_abs",
                     "oracle_csr_softmax", "oracle_csr_softmax_f32", "oracle_csr_softmax_backward",
                     "oracle_csr_transpose"):
            getattr(lib, name).restype = None
        _lib = lib
    return _lib


def _c(a, dt):
    return np.ascontiguousarray(a, dtype=dt)


def spmm(m, k, values, row_offsets, column_indices, dense, bias=None, relu=False):
    """fp64-accumulate SpMM (spmm_test.cu.cc:58-73) + bias/ReLU epilogue (cuda_spmm.cu.cc:267-277).

    ``values`` [nnz] or [R, nnz]; ``dense`` [k, n] or [R, k, n]; a 1-D ``values`` with a 3-D
    ``dense`` is the left_spmm case (values shared).  Returns [m, n] or [R, m, n]."""
    lib = _load()
    values, dense = _c(values, np.float32), _c(dense, np.float32)
    ro, ci = _c(row_offsets, np.int32), _c(column_indices, np.int32)
    if dense.ndim == 2:
        out = np.empty((m, dense.shape[1]), dtype=np.float32)
        b = None if bias is None else _c(bias, np.float32).ctypes.data_as(ctypes.c_void_p)
        lib.oracle_spmm(m, k, dense.shape[1], values.reshape(-1), ro, ci, dense, b, int(relu), out)
        return out
    reps = dense.shape[0]
    outs = [spmm(m, k, values if values.ndim == 1 else values[r], ro, ci, dense[r], bias, relu)
            for r in range(reps)]
    return np.stack(outs)


def spmm_abs(m, k, values, row_offsets, column_indices, dense):
    """sum_l |a_l * b_lj| -- the forward-error scale of each output element."""
    lib = _load()
    dense = _c(dense, np.float32)
    out = np.empty((m, dense.shape[1]), dtype=np.float32)
    lib.oracle_spmm_abs(m, k, dense.shape[1], _c(values, np.float32), _c(row_offsets, np.int32),
                        _c(column_indices, np.int32), dense, out)
    return out


def sddmm(m, n, row_offsets, column_indices, lhs, rhs):
    """fp64 sampled dot products (sddmm_test.cu.cc:59-73).  lhs [m,k]/[R,m,k], rhs [n,k]/[R,n,k]."""
    lib = _load()
    lhs, rhs = _c(lhs, np.float32), _c(rhs, np.float32)
    ro, ci = _c(row_offsets, np.int32), _c(column_indices, np.int32)
    if lhs.ndim == 2:
        out = np.empty(ci.shape[0], dtype=np.float32)
        lib.oracle_sddmm(m, lhs.shape[1], n, ro, ci, lhs, rhs, out)
        return out
    return np.stack([sddmm(m, n, ro, ci, lhs[r], rhs[r]) for r in range(lhs.shape[0])])


def sddmm_abs(m, n, row_offsets, column_indices, lhs, rhs):
    lib = _load()
    lhs, rhs = _c(lhs, np.float32), _c(rhs, np.float32)
    ci = _c(column_indices, np.int32)
    out = np.empty(ci.shape[0], dtype=np.float32)
    lib.oracle_sddmm_abs(m, lhs.shape[1], n, _c(row_offsets, np.int32), ci, lhs, rhs, out)
    return out


def csr_softmax(values, row_offsets, scale: float = 1.0):
    """fp64 evaluation of sparse_softmax.cu.cc:40-68; values [nnz] or [R, nnz]."""
    lib = _load()
    values, ro = _c(values, np.float32), _c(row_offsets, np.int32)
    if values.ndim == 1:
        out = np.zeros_like(values)
        lib.oracle_csr_softmax(ro.shape[0] - 1, values, ro, float(scale), out)
        return out
    return np.stack([csr_softmax(values[r], ro, scale) for r in range(values.shape[0])])


def csr_softmax_f32(values, row_offsets):
    """The kernel's own fp32 lane order (see oracle.c); 1-D values only."""
    lib = _load()
    values, ro = _c(values, np.float32), _c(row_offsets, np.int32)
    out = np.zeros_like(values)
    lib.oracle_csr_softmax_f32(ro.shape[0] - 1, values, ro, out)
    return out


def csr_softmax_backward(y, dy, row_offsets, scale: float = 1.0):
    lib = _load()
    y, dy, ro = _c(y, np.float32), _c(dy, np.float32), _c(row_offsets, np.int32)
    if y.ndim == 1:
        out = np.zeros_like(y)
        lib.oracle_csr_softmax_backward(ro.shape[0] - 1, y, dy, ro, float(scale), out)
        return out
    return np.stack([csr_softmax_backward(y[r], dy[r], ro, scale) for r in range(y.shape[0])])


def csr_transpose(m, n, values, row_offsets, column_indices):
    """Stable CSR transpose.  Returns (values_t, row_offsets_t, column_indices_t, perm);
    values may be [R, nnz] (each replica permuted alike)."""
    lib = _load()
    values = _c(values, np.float32)
    ro, ci = _c(row_offsets, np.int32), _c(column_indices, np.int32)
    nnz = ci.shape[0]
    v1 = values if values.ndim == 1 else _c(values[0], np.float32)
    vt = np.empty(nnz, dtype=np.float32)
    rot = np.empty(n + 1, dtype=np.int32)
    cit = np.empty(nnz, dtype=np.int32)
    perm = np.empty(nnz, dtype=np.int32)
    lib.oracle_csr_transpose(m, n, nnz, v1, ro, ci, vt, rot, cit, perm)
    if values.ndim == 2:
        vt = np.ascontiguousarray(values[:, perm])
    return vt, rot, cit, perm


# ---------------------------------------------------------------------------------------------
# The reference's own CPU-runnable ground truth: dense torch masked matmul.

def dense_spmm(dense_a, rhs):
    """tests/test_spmm.py:9-10,32-34 -- ``torch.matmul(sparse_as_dense, dense)`` (CPU, fp32)."""
    import torch
    return torch.matmul(torch.as_tensor(dense_a), torch.as_tensor(rhs))


def dense_sddmm(lhs, rhs, mask):
    """tests/test_sddmm.py:8-13 / tests/test_sddmm_3d.py:9-14 -- ``lhs @ rhs^T`` then mask."""
    import torch
    lhs, rhs, mask = torch.as_tensor(lhs), torch.as_tensor(rhs), torch.as_tensor(mask)
    return torch.matmul(lhs, rhs.transpose(-2, -1)) * mask


def dense_softmax(dense_scores, mask):
    """tests/test_softmax.py:9-22 -- zeros -> -1e9, ``torch.nn.Softmax(dim=-1)``, re-mask."""
    import torch
    x = torch.as_tensor(dense_scores).clone()
    mask = torch.as_tensor(mask).bool()
    x[~mask] = -1e9
    return torch.nn.Softmax(dim=-1)(x) * mask
