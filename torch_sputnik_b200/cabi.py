"""ctypes binding of the C-ABI (include/tsp.h) -- for tests and harnesses that want to call
libtsp.so directly with raw device pointers, without going through the torch shim."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "lib", "libtsp.so")

# Every symbol include/tsp.h declares.
SYMBOLS = (
    "tsp_version", "tsp_last_error", "tsp_spmm_f32", "tsp_spmm_perm_f32", "tsp_sddmm_f32", "tsp_sddmm_reduce_f32", "tsp_csr_softmax_f32",
    "tsp_csr_softmax_backward_f32", "tsp_csr_transpose_workspace", "tsp_csr_transpose_f32",
    "tsp_permute_values_f32", "tsp_sparse_attention_supported", "tsp_sparse_attention_f32", "tsp_set_variant",
    "tsp_launch_count",
    "tsp_plan_create", "tsp_plan_destroy", "tsp_plan_sorted", "tsp_plan_transpose", "tsp_plan_spmm_f32",
    "tsp_plan_sddmm_f32",
    "tsp_spmm_f16", "tsp_spmm_bf16", "tsp_sddmm_f16", "tsp_sddmm_bf16", "tsp_csr_softmax_f16", "tsp_csr_softmax_bf16",
    "tsp_permute_values_b16", "tsp_plan_spmm_b16", "tsp_plan_sddmm_b16",
    "tsp_plan_permute_values_f32", "tsp_plan_permute_values_b16", "tsp_csr_transpose_noalloc_f32",
    "tsp_sparse_attention_supported_b16", "tsp_sparse_attention_b16",
    "tsp_sddmm_many_f32", "tsp_spmm_many_f32", "tsp_csr_softmax_many_f32", "tsp_permute_values_many_f32",
)

_lib = None


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(f"{LIB_PATH} not built; run `python torch_sputnik_b200/build.py`")
        lib = ctypes.CDLL(LIB_PATH)
        i, vp, f, sz = ctypes.c_int, ctypes.c_void_p, ctypes.c_float, ctypes.c_size_t
        lib.tsp_version.restype = i
        lib.tsp_last_error.restype = ctypes.c_char_p
        lib.tsp_spmm_f32.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, i, vp, vp]
        lib.tsp_spmm_perm_f32.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, i, vp, vp]
        lib.tsp_sddmm_f32.argtypes = [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_sddmm_reduce_f32.argtypes = [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_csr_softmax_f32.argtypes = [i, i, i, f, vp, vp, vp, vp, vp]
        lib.tsp_csr_softmax_backward_f32.argtypes = [i, i, i, f, vp, vp, vp, vp, vp, vp]
        lib.tsp_csr_transpose_workspace.argtypes = [i, i, i]
        lib.tsp_csr_transpose_workspace.restype = sz
        lib.tsp_csr_transpose_f32.argtypes = [i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp, sz, vp]
        lib.tsp_csr_transpose_noalloc_f32.argtypes = [i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_permute_values_f32.argtypes = [i, i, vp, vp, vp, vp]
        lib.tsp_sparse_attention_supported.argtypes = [i, i]
        lib.tsp_sparse_attention_f32.argtypes = [i, i, i, i, i, f, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_sparse_attention_supported_b16.argtypes = [i, i]
        lib.tsp_sparse_attention_b16.argtypes = [i, i, i, i, i, i, f, vp, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_sddmm_many_f32.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_spmm_many_f32.argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, vp]
        lib.tsp_csr_softmax_many_f32.argtypes = [i, i, i, i, vp, f, vp, vp, vp, vp]
        lib.tsp_permute_values_many_f32.argtypes = [i, i, i, vp, vp, vp, vp, vp]
        lib.tsp_set_variant.argtypes = [ctypes.c_char_p, i]
        lib.tsp_launch_count.restype = ctypes.c_ulonglong
        for name in ("tsp_spmm_f16", "tsp_spmm_bf16"):
            getattr(lib, name).argtypes = [i, i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp, i, vp, vp]
        for name in ("tsp_sddmm_f16", "tsp_sddmm_bf16"):
            getattr(lib, name).argtypes = [i, i, i, i, i, vp, vp, vp, vp, vp, vp, vp]
        for name in ("tsp_csr_softmax_f16", "tsp_csr_softmax_bf16"):
            getattr(lib, name).argtypes = [i, i, i, f, vp, vp, vp, vp, vp]
        lib.tsp_permute_values_b16.argtypes = [i, i, vp, vp, vp, vp]
        pp = ctypes.POINTER(vp)
        lib.tsp_plan_create.argtypes = [i, i, i, vp, vp, vp, pp]
        lib.tsp_plan_destroy.argtypes = [vp]
        lib.tsp_plan_destroy.restype = None
        lib.tsp_plan_sorted.argtypes = [vp]
        lib.tsp_plan_transpose.argtypes = [vp, vp, pp, pp, pp, pp]
        lib.tsp_plan_spmm_f32.argtypes = [vp, i, i, i, vp, vp, vp, vp, vp, i, vp, vp]
        lib.tsp_plan_sddmm_f32.argtypes = [vp, i, i, i, vp, vp, vp, vp, vp]
        lib.tsp_plan_spmm_b16.argtypes = [vp, i, i, i, i, vp, vp, vp, vp, vp, i, vp, vp]
        lib.tsp_plan_sddmm_b16.argtypes = [vp, i, i, i, vp, vp, vp, vp, vp]
        lib.tsp_plan_permute_values_f32.argtypes = [vp, i, vp, vp, vp]
        lib.tsp_plan_permute_values_b16.argtypes = [vp, i, vp, vp, vp]
        _lib = lib
    return _lib


def last_error() -> str:
    return load().tsp_last_error().decode()
