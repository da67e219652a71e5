"""torch_sputnik_b200 -- B200-native CSR sparse kernels behind the ``torch_sputnik`` operator API.

Importing this package loads the in-tree extension module and registers it as
``sys.modules['torch_sputnik']``, so code written against the reference
(``import torch_sputnik``; modules/spmm.py, sddmm.py, sparse_linear.py, sparse_attention.py of
mabdullahsoyturk/Torch-Sputnik) runs unchanged.  There is no CPU or library fallback: if the
native artefacts are missing the import raises, with the build command in the message.
"""
from __future__ import annotations

import importlib.util
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))


def _load_extension():
    import sysconfig

    import torch  # noqa: F401  (libtorch / libc10 / libcudart must be loaded first)

    suffix = sysconfig.get_config_var("EXT_SUFFIX") or ".so"
    path = os.path.join(_HERE, "torch_sputnik" + suffix)
    lib = os.path.join(_HERE, "lib", "libtsp.so")
    if not (os.path.exists(path) and os.path.exists(lib)):
        raise ImportError(
            "torch_sputnik_b200: native extension not built (expected %s and %s). "
            "Run `python torch_sputnik_b200/build.py` (needs nvcc, sm_100a). "
            "There is no CPU fallback." % (path, lib))
    spec = importlib.util.spec_from_file_location("torch_sputnik", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


torch_sputnik = _load_extension()
sys.modules.setdefault("torch_sputnik", torch_sputnik)

spmm = torch_sputnik.spmm
left_spmm = torch_sputnik.left_spmm
sddmm = torch_sputnik.sddmm
sparse_softmax = torch_sputnik.sparse_softmax
csr_transpose = torch_sputnik.csr_transpose
spmm_bias_relu = torch_sputnik.spmm_bias_relu
spmm_bias = torch_sputnik.spmm_bias
sddmm_reduce = torch_sputnik.sddmm_reduce
sparse_softmax_scaled = torch_sputnik.sparse_softmax_scaled
sparse_softmax_backward = torch_sputnik.sparse_softmax_backward
csr_transpose_with_perm = torch_sputnik.csr_transpose_with_perm
permute_values = torch_sputnik.permute_values
sparse_attention = torch_sputnik.sparse_attention
spmm_perm = torch_sputnik.spmm_perm
spmm_many_mask = torch_sputnik.spmm_many_mask
sddmm_many_mask = torch_sputnik.sddmm_many_mask
sparse_softmax_many_mask = torch_sputnik.sparse_softmax_many_mask
csr_transpose_many_mask = torch_sputnik.csr_transpose_many_mask
sparse_softmax_many_mask_scaled = torch_sputnik.sparse_softmax_many_mask_scaled
transposed_topology = torch_sputnik.transposed_topology
plan_policy = torch_sputnik.plan_policy
plan_cache_clear = torch_sputnik.plan_cache_clear
plan_cache_stats = torch_sputnik.plan_cache_stats
set_variant = torch_sputnik.set_variant
launch_count = torch_sputnik.launch_count

__all__ = [
    "torch_sputnik", "spmm", "left_spmm", "sddmm", "sparse_softmax", "csr_transpose",
    "spmm_bias_relu", "spmm_bias", "sddmm_reduce", "sparse_softmax_scaled", "sparse_softmax_backward",
    "csr_transpose_with_perm", "permute_values", "sparse_attention", "spmm_perm", "spmm_many_mask", "sddmm_many_mask",
    "sparse_softmax_many_mask", "csr_transpose_many_mask", "sparse_softmax_many_mask_scaled", "transposed_topology",
    "plan_policy", "plan_cache_clear", "plan_cache_stats", "set_variant", "launch_count",
]
