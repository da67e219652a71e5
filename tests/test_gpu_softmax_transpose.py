"""GPU parity: CSR softmax (fwd/bwd) within 1e-5 relative of the fp64 oracle; csr_transpose
bit-exact in offsets, indices and value placement."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from helpers import to_dev
from oracle.csr import dense_operand, random_csr, row_of

pytestmark = pytest.mark.gpu


# ----------------------------------------------------------------------------- softmax
@pytest.mark.parametrize("m,n,density", [(72, 72, 0.1), (1, 5, 1.0), (64, 64, 0.0), (200, 33, 0.5),
                                         (128, 4096, 0.1), (16, 4096, 1.0), (300, 700, 0.002),
                                         (4096, 4096, 0.1)])
def test_softmax_vs_oracle(ops, dev, m, n, density):
    a = random_csr(m, n, density, seed=m + n, values="pm1")
    a.values = (a.values * 8).astype(np.float32)            # scores with a real dynamic range
    v, ri, ro, ci = to_dev(a, dev)
    got = ops.sparse_softmax(v, ri, ro, ci).cpu().numpy()
    ref = oracle.csr_softmax(a.values, a.row_offsets)
    assert got.shape == ref.shape
    if a.nnz:
        np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-30)
        # the reference test's ground truth: dense torch softmax with zeros -> -1e9 (tests/test_softmax.py:9-22)
        if m * n <= 1 << 20:
            dense = a.to_dense()
            want = oracle.dense_softmax(dense, dense != 0).numpy()[row_of(a.row_offsets), a.column_indices]
            np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-12)


def test_softmax_replicated_scaled_and_backward(ops, dev):
    R, s = 6, 256
    a = random_csr(s, s, 0.1, seed=1, values="pm1")
    x = dense_operand((R, a.nnz), 2, "pm1") * 6
    _, ri, ro, ci = to_dev(a, dev)
    xd = torch.from_numpy(x).to(dev)
    y = ops.sparse_softmax(xd, ri, ro, ci)
    assert tuple(y.shape) == (R, a.nnz)
    np.testing.assert_allclose(y.cpu().numpy(), oracle.csr_softmax(x, a.row_offsets), rtol=1e-5)
    ys = ops.sparse_softmax_scaled(xd, ri, ro, ci, 0.125)
    np.testing.assert_allclose(ys.cpu().numpy(), oracle.csr_softmax(x, a.row_offsets, 0.125), rtol=1e-5)
    dy = dense_operand((R, a.nnz), 3, "pm1")
    dx = ops.sparse_softmax_backward(ys, torch.from_numpy(dy).to(dev), ri, ro, ci, 0.125).cpu().numpy()
    ref = oracle.csr_softmax_backward(ys.cpu().numpy(), dy, a.row_offsets, 0.125)
    np.testing.assert_allclose(dx, ref, rtol=1e-4, atol=1e-7)


def test_softmax_extreme_values_and_long_rows(ops, dev):
    """Rows longer than the register path (32*REG), +-huge scores, single-element rows."""
    ro = np.array([0, 1, 1, 3001, 3004], dtype=np.int32)
    x = np.concatenate([[1e30], np.linspace(-80, 80, 3000), [-1e30, 0.0, 1e30]]).astype(np.float32)
    m = 4
    t = lambda a: torch.from_numpy(a).to(dev)
    got = ops.sparse_softmax(t(x), t(np.arange(m, dtype=np.int32)), t(ro), t(np.zeros(len(x), np.int32))).cpu().numpy()
    ref = oracle.csr_softmax(x, ro)
    assert got[0] == 1.0
    np.testing.assert_allclose(got, ref, rtol=1e-5, atol=1e-37)
    assert np.isfinite(got).all()


# --------------------------------------------------------------------------- transpose
@pytest.mark.parametrize("m,n,density", [(4, 4, 0.6), (72, 64, 0.1), (64, 72, 1.0), (300, 1000, 0.02),
                                         (1000, 37, 0.3), (5, 70000, 0.001), (70000, 3, 0.4),
                                         (2, 2, 0.0), (4096, 4096, 0.1), (4096, 4096, 0.5), (513, 300000, 0.01)])
def test_transpose_bit_exact(ops, dev, m, n, density):
    a = random_csr(m, n, density, seed=m + 2 * n, values="pm1")
    v, _, ro, ci = to_dev(a, dev)
    vt, rot, cit = ops.csr_transpose(m, n, v, ro, ci)
    evt, erot, ecit, eperm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    assert rot.dtype == torch.int32 and cit.dtype == torch.int32 and vt.dtype == torch.float32
    np.testing.assert_array_equal(rot.cpu().numpy(), erot)
    np.testing.assert_array_equal(cit.cpu().numpy(), ecit)
    np.testing.assert_array_equal(vt.cpu().numpy(), evt)
    vt2, rot2, cit2, perm = ops.csr_transpose_with_perm(m, n, v, ro, ci)
    np.testing.assert_array_equal(perm.cpu().numpy(), eperm)
    assert torch.equal(vt2, vt) and torch.equal(rot2, rot) and torch.equal(cit2, cit)
    if a.nnz:
        assert torch.equal(ops.permute_values(v, perm), vt)


@pytest.mark.parametrize("m,n,density", [(3072, 768, 0.1), (768, 3072, 0.1), (1, 5000, 0.5), (9000, 48000, 0.0005), (2000, 30000, 0.05), (800, 49152, 0.15),
                                         (2500, 1, 0.7), (257, 1025, 0.33)])
def test_transpose_counting_and_radix_paths_agree(ops, dev, m, n, density):
    """Both sorts behind csr_transpose (3-launch counting path, multi-pass radix path) against the oracle;
    3072 x 768 at 10 % is the sparse linear layer of tests/test_left_spmm.py:101-154."""
    a = random_csr(m, n, density, seed=7 * m + n, values="pm1")
    v, _, ro, ci = to_dev(a, dev)
    evt, erot, ecit, eperm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    ops.plan_policy(0)
    try:
        for variant in (0, 1):
            ops.set_variant("transpose", variant)
            vt, rot, cit, perm = ops.csr_transpose_with_perm(m, n, v, ro, ci)
            np.testing.assert_array_equal(rot.cpu().numpy(), erot)
            np.testing.assert_array_equal(cit.cpu().numpy(), ecit)
            np.testing.assert_array_equal(vt.cpu().numpy(), evt)
            np.testing.assert_array_equal(perm.cpu().numpy(), eperm)
    finally:
        ops.set_variant("transpose", 0)
        ops.plan_policy(1)


def test_transpose_known_answer_4x4_empty_first_row(ops, dev):
    """tests/test_transpose.py:27-32."""
    t = lambda a, dt: torch.tensor(a, dtype=dt, device=dev)
    vt, rot, cit = ops.csr_transpose(4, 4, t([1, 2, 3, 4, 5, 6, 7], torch.float32),
                                     t([0, 0, 2, 4, 7], torch.int32), t([0, 2, 1, 3, 0, 1, 3], torch.int32))
    assert rot.tolist() == [0, 2, 4, 5, 7]
    assert cit.tolist() == [1, 3, 2, 3, 1, 2, 3]
    assert vt.tolist() == [1, 5, 3, 6, 2, 4, 7]


def test_transpose_replicated_values_and_involution(ops, dev):
    """[R, nnz] values (what modules/spmm.py:59 passes in 3-D) and transpose(transpose(A)) == A."""
    m, n = 333, 444
    a = random_csr(m, n, 0.07, seed=3, values="pm1")
    vals = np.stack([a.values, -a.values, a.values * 3]).astype(np.float32)
    _, _, ro, ci = to_dev(a, dev)
    vt, rot, cit = ops.csr_transpose(m, n, torch.from_numpy(vals).to(dev), ro, ci)
    evt, erot, ecit, _ = oracle.csr_transpose(m, n, vals, a.row_offsets, a.column_indices)
    np.testing.assert_array_equal(vt.cpu().numpy(), evt)
    np.testing.assert_array_equal(rot.cpu().numpy(), erot)
    vtt, rott, citt = ops.csr_transpose(n, m, vt, rot, cit)
    np.testing.assert_array_equal(vtt.cpu().numpy(), vals)
    np.testing.assert_array_equal(rott.cpu().numpy(), a.row_offsets)
    np.testing.assert_array_equal(citt.cpu().numpy(), a.column_indices)


def test_transpose_duplicate_columns_stay_stable(ops, dev):
    """Duplicates inside a row are legal CSR for cuSPARSE; stability keeps their order."""
    ro = np.array([0, 4, 4, 7], dtype=np.int32)
    ci = np.array([2, 2, 0, 2, 1, 2, 2], dtype=np.int32)
    va = np.arange(1, 8, dtype=np.float32)
    t = lambda x: torch.from_numpy(x).to(dev)
    vt, rot, cit = ops.csr_transpose(3, 3, t(va), t(ro), t(ci))
    evt, erot, ecit, _ = oracle.csr_transpose(3, 3, va, ro, ci)
    np.testing.assert_array_equal(vt.cpu().numpy(), evt)
    np.testing.assert_array_equal(rot.cpu().numpy(), erot)
    np.testing.assert_array_equal(cit.cpu().numpy(), ecit)


def test_transpose_through_raw_c_abi(dev):
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    m, n = 700, 900
    a = random_csr(m, n, 0.05, seed=4, values="pm1")
    v, _, ro, ci = to_dev(a, dev)
    size = lib.tsp_csr_transpose_workspace(m, n, a.nnz)
    ws = torch.empty(size, dtype=torch.uint8, device=dev)
    vt = torch.empty_like(v)
    rot = torch.empty(n + 1, dtype=torch.int32, device=dev)
    cit = torch.empty(a.nnz, dtype=torch.int32, device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    rc = lib.tsp_csr_transpose_f32(m, n, a.nnz, 1, p(v), p(ro), p(ci), p(vt), p(rot), p(cit), None, p(ws), size, stream)
    assert rc == 0, cabi.last_error()
    evt, erot, ecit, _ = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    np.testing.assert_array_equal(vt.cpu().numpy(), evt)
    np.testing.assert_array_equal(rot.cpu().numpy(), erot)
    np.testing.assert_array_equal(cit.cpu().numpy(), ecit)
    rc = lib.tsp_csr_transpose_f32(m, n, a.nnz, 1, p(v), p(ro), p(ci), p(vt), p(rot), p(cit), None, p(ws), 16, stream)
    assert rc == 3 and "workspace" in cabi.last_error()
