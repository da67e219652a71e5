"""GPU: tsp_csr_transpose_noalloc_f32 -- the entry point that allocates nothing (its scratch is a static device
array inside libtsp.so) -- through the raw C-ABI against tsp_csr_transpose_f32 with a caller-supplied workspace.
Integer outputs and moved values: bit-exact, as everywhere else for csr_transpose.  Collected last on purpose."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from oracle.csr import random_csr

pytestmark = pytest.mark.gpu

NOALLOC_SCRATCH_BYTES = 8 << 20   # TSP_NOALLOC_SCRATCH_BYTES, include/tsp.h


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _run_both(lib, cabi, dev, m, n, density, replicas, seed):
    a = random_csr(m, n, density, seed=seed, values="pm1")
    nnz = int(a.nnz)
    g = torch.Generator(device="cpu").manual_seed(seed)
    v = torch.rand((replicas, nnz), generator=g).to(dev)
    ro = torch.from_numpy(a.row_offsets).to(dev)
    ci = torch.from_numpy(a.column_indices).to(dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)

    def outputs():   # filled with a pattern, so that "not written" cannot pass for "equal"
        return (torch.full((replicas, nnz), -7.0, device=dev), torch.full((n + 1,), -7, dtype=torch.int32, device=dev),
                torch.full((nnz,), -7, dtype=torch.int32, device=dev), torch.full((nnz,), -7, dtype=torch.int32, device=dev))

    size = lib.tsp_csr_transpose_workspace(m, n, nnz)
    assert size <= NOALLOC_SCRATCH_BYTES
    ws = torch.empty(size, dtype=torch.uint8, device=dev)
    vt0, rot0, cit0, perm0 = outputs()
    rc = lib.tsp_csr_transpose_f32(m, n, nnz, replicas, _p(v), _p(ro), _p(ci), _p(vt0), _p(rot0), _p(cit0), _p(perm0),
                                   _p(ws), size, stream)
    assert rc == 0, cabi.last_error()
    vt1, rot1, cit1, perm1 = outputs()
    torch.cuda.synchronize()
    rc = lib.tsp_csr_transpose_noalloc_f32(m, n, nnz, replicas, _p(v), _p(ro), _p(ci), _p(vt1), _p(rot1), _p(cit1),
                                           _p(perm1), stream)
    assert rc == 0, cabi.last_error()
    torch.cuda.synchronize()
    assert torch.equal(rot1, rot0) and torch.equal(cit1, cit0) and torch.equal(perm1, perm0) and torch.equal(vt1, vt0)
    # and both against the CPU oracle
    _, erot, ecit, eperm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    np.testing.assert_array_equal(rot1.cpu().numpy(), erot)
    np.testing.assert_array_equal(cit1.cpu().numpy(), ecit)
    np.testing.assert_array_equal(perm1.cpu().numpy(), eperm)
    np.testing.assert_array_equal(vt1.cpu().numpy(), v.cpu().numpy()[:, eperm])


def test_noalloc_transpose_equals_the_workspace_entry_point_twice_in_a_row():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    dev = torch.device("cuda:0")
    torch.cuda.set_device(dev)
    # two calls in a row on the shared static scratch: a larger problem, then a smaller one of another shape (the
    # second sees the first one's leftovers), then the first again
    _run_both(lib, cabi, dev, 3072, 768, 0.1, 2, seed=11)      # sparse linear weight, tests/test_left_spmm.py:101-154
    _run_both(lib, cabi, dev, 300, 1000, 0.02, 1, seed=12)
    _run_both(lib, cabi, dev, 3072, 768, 0.1, 2, seed=11)
    _run_both(lib, cabi, dev, 512, 60000, 0.002, 1, seed=13)   # wide and nearly empty: the radix path's buffers


def test_noalloc_transpose_rejects_problems_beyond_the_static_scratch():
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device")
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    m = n = 8192
    nnz = 1 << 25
    assert lib.tsp_csr_transpose_workspace(m, n, nnz) > NOALLOC_SCRATCH_BYTES
    rc = lib.tsp_csr_transpose_noalloc_f32(m, n, nnz, 0, None, None, None, None, None, None, None, None)
    assert rc == 3
    msg = cabi.last_error()
    assert "static" in msg and str(NOALLOC_SCRATCH_BYTES) in msg
    assert lib.tsp_csr_transpose_noalloc_f32(-1, 4, 0, 0, None, None, None, None, None, None, None, None) == 1
