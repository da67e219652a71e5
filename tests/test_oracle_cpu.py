"""CPU: the oracle against independent restatements (numpy fp64, torch dense, torch CSR) and
against hand-computed known answers.  These pin the oracle's OWN correctness; the pin to the
reference's kernels is tests/test_oracle_golden.py."""
import numpy as np
import pytest
import torch

import oracle
from oracle.csr import Csr, dense_operand, random_csr, row_of


SPMM_PROBLEMS = [  # third_party/sputnik/sputnik/spmm/spmm_test.cu.cc:80-95 (CPU-sized subset)
    (4, 32, 32, 32), (4, 32, 48, 128), (7, 96, 40, 333), (11, 55, 8, 293), (4, 128, 128, 512),
    (64, 512, 512, 16384),
]
SDDMM_PROBLEMS = [  # third_party/sputnik/sputnik/sddmm/sddmm_test.cu.cc:76-91 (m, k, n, nnz)
    (1, 32, 32, 32), (1, 64, 32, 32), (1, 32, 64, 64), (1, 32, 64, 32), (1, 32, 64, 16),
    (2, 32, 32, 64), (1, 28, 32, 32), (1024, 80, 1024, 131072),
]


@pytest.mark.parametrize("m,k,n,nnz", SPMM_PROBLEMS)
def test_spmm_oracle_vs_numpy_fp64(m, k, n, nnz):
    a = random_csr(m, k, nnz=nnz, seed=m + n, values="pm1")
    b = dense_operand((k, n), 7, "pm1")
    ref = (a.to_dense().astype(np.float64) @ b.astype(np.float64))
    got = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, b)
    np.testing.assert_allclose(got, ref.astype(np.float32), rtol=0, atol=1e-6)
    # the reference tests' own check: dense torch matmul within 1e-2 (tests/test_spmm.py:36)
    dense = oracle.dense_spmm(a.to_dense(), b).numpy()
    assert np.all(np.abs(got - dense) < 1e-2)


def test_spmm_bias_relu_oracle():
    a = random_csr(9, 16, 0.4, seed=3, values="pm1")
    b = dense_operand((16, 12), 4, "pm1")
    bias = np.linspace(-1, 1, 9).astype(np.float32)
    got = oracle.spmm(9, 16, a.values, a.row_offsets, a.column_indices, b, bias=bias, relu=True)
    ref = np.maximum(a.to_dense().astype(np.float64) @ b + bias[:, None], 0)
    np.testing.assert_allclose(got, ref, atol=1e-6)
    # tests/test_spmm_bias_relu.py:27-43: full mask, bias of ones, positive inputs -> dense + 1
    a = random_csr(8, 8, 1.0, seed=1, values="u01")
    b = dense_operand((8, 8), 2, "u01")
    got = oracle.spmm(8, 8, a.values, a.row_offsets, a.column_indices, b, bias=np.ones(8, np.float32),
                      relu=True)
    np.testing.assert_allclose(got, a.to_dense() @ b + 1, atol=1e-5)


@pytest.mark.parametrize("m,k,n,nnz", SDDMM_PROBLEMS)
def test_sddmm_oracle_vs_numpy_fp64(m, k, n, nnz):
    a = random_csr(m, n, nnz=nnz, seed=k, values="pm1")
    lhs = dense_operand((m, k), 1, "pm1")
    rhs = dense_operand((n, k), 2, "pm1")
    full = lhs.astype(np.float64) @ rhs.astype(np.float64).T
    ref = full[row_of(a.row_offsets), a.column_indices]
    got = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    np.testing.assert_allclose(got, ref.astype(np.float32), rtol=0, atol=1e-6)
    mask = (a.to_dense() != 0).astype(np.float32)
    dense = oracle.dense_sddmm(lhs, rhs, mask).numpy()
    assert np.all(np.abs(dense[row_of(a.row_offsets), a.column_indices] - got) < 1e-4 * max(1, k / 32))


def test_softmax_oracle_vs_torch_dense():
    a = random_csr(72, 72, 0.1, seed=5, values="pm1")   # tests/test_softmax.py:25-33 shape
    y = oracle.csr_softmax(a.values, a.row_offsets)
    dense = a.to_dense()
    mask = dense != 0
    ref = oracle.dense_softmax(dense, mask).numpy()
    np.testing.assert_allclose(y, ref[row_of(a.row_offsets), a.column_indices], rtol=2e-6, atol=1e-8)
    y32 = oracle.csr_softmax_f32(a.values, a.row_offsets)
    np.testing.assert_allclose(y32, y, rtol=1e-5, atol=1e-8)
    # rows sum to one, empty rows untouched
    sums = np.add.reduceat(np.append(y, 0), a.row_offsets[:-1].clip(max=len(y)))[np.diff(a.row_offsets) > 0]
    np.testing.assert_allclose(sums, 1.0, rtol=1e-5)
    ys = oracle.csr_softmax(a.values, a.row_offsets, scale=0.125)
    np.testing.assert_allclose(ys, oracle.csr_softmax(a.values * np.float32(0.125), a.row_offsets), rtol=1e-6)


def test_softmax_backward_oracle_vs_autograd():
    a = random_csr(16, 40, 0.3, seed=6, values="pm1")
    x = torch.tensor(a.to_dense(), dtype=torch.float64)
    mask = torch.tensor(a.to_dense() != 0)
    xs = (x * 0.25).masked_fill(~mask, -float("inf")).requires_grad_(False)
    xv = torch.tensor(a.values, dtype=torch.float64, requires_grad=True)
    dense = torch.full((16, 40), -float("inf"), dtype=torch.float64)
    r = torch.tensor(row_of(a.row_offsets), dtype=torch.long)
    c = torch.tensor(a.column_indices, dtype=torch.long)
    dense = dense.index_put((r, c), xv * 0.25)
    rows_with = torch.tensor(np.diff(a.row_offsets) > 0)
    y = torch.softmax(dense[rows_with], dim=-1)
    g = torch.tensor(dense_operand((16, 40), 9, "pm1"), dtype=torch.float64)
    (y * g[rows_with]).sum().backward()
    yv = oracle.csr_softmax(a.values, a.row_offsets, 0.25)
    dy = g.numpy()[r.numpy(), c.numpy()].astype(np.float32)
    dx = oracle.csr_softmax_backward(yv, dy, a.row_offsets, 0.25)
    np.testing.assert_allclose(dx, xv.grad.numpy(), rtol=1e-4, atol=1e-7)
    del xs


def test_transpose_known_answer_4x4_empty_first_row():
    """tests/test_transpose.py:27-32: a 4x4 matrix whose first row is zero."""
    dense = np.array([[0, 0, 0, 0], [1, 0, 2, 0], [0, 3, 0, 4], [5, 6, 0, 7]], dtype=np.float32)
    ro = np.array([0, 0, 2, 4, 7], dtype=np.int32)
    ci = np.array([0, 2, 1, 3, 0, 1, 3], dtype=np.int32)
    va = np.array([1, 2, 3, 4, 5, 6, 7], dtype=np.float32)
    vt, rot, cit, perm = oracle.csr_transpose(4, 4, va, ro, ci)
    assert rot.tolist() == [0, 2, 4, 5, 7]
    assert cit.tolist() == [1, 3, 2, 3, 1, 2, 3]
    assert vt.tolist() == [1, 5, 3, 6, 2, 4, 7]
    assert perm.tolist() == [0, 4, 2, 5, 1, 3, 6]
    t = Csr(4, 4, vt, np.arange(4, dtype=np.int32), rot, cit).to_dense()
    np.testing.assert_array_equal(t, dense.T)


@pytest.mark.parametrize("m,n,density", [(4, 4, 0.5), (72, 64, 0.1), (300, 1000, 0.02), (1000, 37, 0.3),
                                         (64, 64, 1.0), (5, 70000, 0.001)])
def test_transpose_oracle_vs_numpy_and_torch(m, n, density):
    a = random_csr(m, n, density, seed=m * 3 + n, values="pm1")
    vt, rot, cit, perm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    order = np.argsort(a.column_indices, kind="stable")
    np.testing.assert_array_equal(perm, order.astype(np.int32))
    np.testing.assert_array_equal(vt, a.values[order])
    np.testing.assert_array_equal(cit, row_of(a.row_offsets)[order])
    np.testing.assert_array_equal(rot, np.concatenate([[0], np.cumsum(np.bincount(a.column_indices, minlength=n))]))
    # torch's own CSR of the dense transpose (what a caller would compare with)
    csr = torch.tensor(a.to_dense().T.copy()).to_sparse_csr()
    np.testing.assert_array_equal(rot, csr.crow_indices().numpy().astype(np.int32))
    np.testing.assert_array_equal(cit, csr.col_indices().numpy().astype(np.int32))
    np.testing.assert_array_equal(vt, csr.values().numpy())
    # replicated values share the permutation
    v2 = np.stack([a.values, a.values * 2])
    vt2, *_ = oracle.csr_transpose(m, n, v2, a.row_offsets, a.column_indices)
    np.testing.assert_array_equal(vt2[1], vt * 2)


def test_generator_semantics():
    """tests/connectors.py:47-59: exact nnz, rounded up to round_to; sorted columns; swizzles."""
    a = random_csr(72, 64, 0.1, seed=0, round_to=4)
    assert a.nnz == 464 and a.nnz % 4 == 0            # round(0.9*4608)=4147 dormant -> 461 -> 464
    for i in range(72):
        seg = a.column_indices[a.row_offsets[i]:a.row_offsets[i + 1]]
        assert np.all(np.diff(seg) > 0)
    lens = np.diff(a.row_offsets)
    assert np.all(np.diff(lens[a.row_indices]) <= 0)   # heaviest first (tests/sparse_matrix.py:22)
    asc = random_csr(72, 64, 0.1, seed=0, round_to=4, swizzle="ascending")
    assert np.all(np.diff(lens[asc.row_indices]) >= 0)
    assert sorted(a.row_indices.tolist()) == list(range(72))
    b = random_csr(72, 64, 0.1, seed=0, round_to=4)
    np.testing.assert_array_equal(a.column_indices, b.column_indices)
    assert random_csr(8, 8, 0.0, seed=1).nnz == 0
    assert random_csr(8, 8, 1.0, seed=1).nnz == 64


def test_attention_chain_oracle_vs_dense_masked_attention():
    """The checker of tests/test_gpu_attention.py (oracle sddmm -> softmax(scale) -> spmm) against
    dense masked softmax attention in fp64 (modules/sparse_attention.py:66-82 semantics)."""
    import torch
    from oracle import cpu as ocpu
    from oracle.csr import dense_operand, from_dense_mask
    rng = np.random.default_rng(3)
    m, n, d, scale = 40, 56, 32, 0.2
    mask = rng.random((m, n)) < 0.3
    mask[:, 0] = True                                   # no empty row: the dense softmax needs one entry
    a = from_dense_mask(mask, seed=1)
    q, k, v = dense_operand((m, d), 1, "pm1"), dense_operand((n, d), 2, "pm1"), dense_operand((n, d), 3, "pm1")
    s = ocpu.sddmm(m, n, a.row_offsets, a.column_indices, q, k)
    p = ocpu.csr_softmax(s, a.row_offsets, scale)
    out = ocpu.spmm(m, n, p, a.row_offsets, a.column_indices, v)
    scores = torch.from_numpy(q).double() @ torch.from_numpy(k).double().T * scale
    scores = scores.masked_fill(~torch.from_numpy(mask), float("-inf"))
    ref = torch.softmax(scores, dim=-1) @ torch.from_numpy(v).double()
    np.testing.assert_allclose(out, ref.numpy(), rtol=2e-6, atol=2e-6)
