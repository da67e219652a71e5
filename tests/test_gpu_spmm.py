"""GPU parity: CUDA SpMM (through the torch_sputnik shim and the raw C-ABI) vs the oracle."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from helpers import ABS_SPUTNIK, ABS_TORCH, assert_close_fwd, spmm_case, to_dev
from oracle.csr import dense_operand, random_csr

pytestmark = pytest.mark.gpu

# third_party/sputnik/sputnik/spmm/spmm_test.cu.cc:80-95  (m, k, n, nnz)
SPUTNIK_SPMM_PROBLEMS = [
    (4, 32, 32, 32), (4, 32, 48, 128), (7, 96, 40, 333), (11, 55, 8, 293), (4, 128, 128, 512),
    (64, 512, 512, 16384), (1024, 1024, 512, 131072), (1024, 1024, 256, 131072),
    (1024, 1024, 128, 131072), (2048, 2048, 128, 1024 * 1024),
    (2048 * 3, 2048, 128, 1024 * 1024 * 3), (2048 * 4, 2048, 128, 1024 * 1024 * 4),
]


def run_spmm(ops, dev, a, b, **kw):
    v, ri, ro, ci = to_dev(a, dev)
    return ops.spmm(a.m, a.k, v, ri, ro, ci, torch.from_numpy(b).to(dev), **kw).cpu().numpy()


def check_spmm(ops, dev, a, b, what):
    got = run_spmm(ops, dev, a, b)
    ref = oracle.spmm(a.m, a.k, a.values, a.row_offsets, a.column_indices, b)
    scale = oracle.spmm_abs(a.m, a.k, a.values, a.row_offsets, a.column_indices, b)
    assert got.shape == ref.shape
    assert_close_fwd(got, ref, scale, what=what)
    return got, ref


@pytest.mark.parametrize("m,k,n,nnz", SPUTNIK_SPMM_PROBLEMS)
def test_spmm_sputnik_problem_list(ops, dev, m, k, n, nnz):
    a, b = spmm_case(m, k, n, nnz=nnz, seed=n + m, values="pm1")
    got, ref = check_spmm(ops, dev, a, b, f"spmm {m}x{k}x{n}/{nnz}")
    assert np.max(np.abs(got - ref)) < ABS_SPUTNIK   # NanSensitiveFloatNear(1e-4), spmm_test.cu.cc:151


def test_spmm_reference_test_shape(ops, dev):
    """tests/test_spmm.py:13-40: 72x64x72, sparsity 0.9, round_to=4, U[0,1), |diff| < 1e-2 vs dense."""
    a, b = spmm_case(72, 64, 72, density=0.1, seed=0, values="u01", round_to=4)
    got, _ = check_spmm(ops, dev, a, b, "72x64x72")
    dense = oracle.dense_spmm(a.to_dense(), b).numpy()
    assert int((np.abs(got - dense) < ABS_TORCH).sum()) == 72 * 72


def test_spmm_baseline_config_64(ops, dev):
    """BASELINE.json configs[0]: M=N=K=64, density 0.25."""
    a, b = spmm_case(64, 64, 64, density=0.25, seed=1, values="u01")
    assert a.nnz == 1024
    got, _ = check_spmm(ops, dev, a, b, "64^3")
    dense = oracle.dense_spmm(a.to_dense(), b).numpy()
    assert np.all(np.abs(got - dense) < ABS_TORCH)


@pytest.mark.parametrize("n", [1, 2, 3, 7, 8, 31, 33, 64, 65, 72, 100, 128, 130, 192, 250, 256, 260, 516])
def test_spmm_ragged_n(ops, dev, n):
    a, b = spmm_case(37, 53, n, density=0.3, seed=n, values="pm1")
    check_spmm(ops, dev, a, b, f"n={n}")


@pytest.mark.parametrize("density", [0.0, 0.01, 0.5, 1.0])
def test_spmm_density_extremes_and_empty_rows(ops, dev, density):
    a, b = spmm_case(129, 96, 136, density=density, seed=3, values="pm1")
    got, _ = check_spmm(ops, dev, a, b, f"density={density}")
    empty = np.diff(a.row_offsets) == 0
    assert np.all(got[empty] == 0.0)            # every output element is written, zeros included


def test_spmm_row_swizzle_is_only_a_hint(ops, dev):
    """Results are bit-identical for any permutation in row_indices."""
    a, b = spmm_case(300, 200, 256, density=0.1, seed=4, values="pm1")
    base = run_spmm(ops, dev, a, b)
    rng = np.random.default_rng(0)
    for ri in (np.arange(300), np.argsort(np.diff(a.row_offsets), kind="stable"), rng.permutation(300)):
        a.row_indices = ri.astype(np.int32)
        np.testing.assert_array_equal(run_spmm(ops, dev, a, b), base)
    # int64 row_indices (modules/sparse_linear.py:5-7 returns int64) are accepted
    v, _, ro, ci = to_dev(a, dev)
    got = ops.spmm(300, 200, v, torch.arange(300, device=dev), ro, ci, torch.from_numpy(b).to(dev))
    np.testing.assert_array_equal(got.cpu().numpy(), base)


def test_spmm_unsorted_columns(ops, dev):
    """The API never promised ascending columns inside a row; shuffle them."""
    a, b = spmm_case(128, 512, 256, density=0.2, seed=5, values="pm1")
    rng = np.random.default_rng(1)
    for i in range(a.m):
        s, e = a.row_offsets[i], a.row_offsets[i + 1]
        p = rng.permutation(e - s)
        a.column_indices[s:e] = a.column_indices[s:e][p]
        a.values[s:e] = a.values[s:e][p]
    check_spmm(ops, dev, a, b, "unsorted columns")


def test_spmm_replicated_3d_and_quirks(ops, dev):
    """src/spmm_cuda.cu:29-30,46: [R,m,n] for R > 1, but 2-D when the replica count is 1."""
    R, m, k, n = 5, 72, 64, 72
    a = random_csr(m, k, 0.1, seed=6, values="u01", round_to=4)
    vals = np.stack([a.values * (r + 1) for r in range(R)]).astype(np.float32)
    dense = dense_operand((R, k, n), 7, "u01")
    v, ri, ro, ci = to_dev(a, dev)
    got = ops.spmm(m, k, torch.from_numpy(vals).to(dev), ri, ro, ci, torch.from_numpy(dense).to(dev))
    assert tuple(got.shape) == (R, m, n)
    ref = oracle.spmm(m, k, vals, a.row_offsets, a.column_indices, dense)
    for r in range(R):
        scale = oracle.spmm_abs(m, k, vals[r], a.row_offsets, a.column_indices, dense[r])
        assert_close_fwd(got[r].cpu().numpy(), ref[r], scale, what=f"replica {r}")
    one = ops.spmm(m, k, torch.from_numpy(vals[:1]).to(dev), ri, ro, ci, torch.from_numpy(dense[:1]).to(dev))
    assert tuple(one.shape) == (m, n)
    np.testing.assert_array_equal(one.cpu().numpy(), got[0].cpu().numpy())


def test_left_spmm_shared_values_always_3d(ops, dev):
    """src/left_replicated_spmm.cu:30-41: one values vector for every replica, output always 3-D."""
    R, m, k, n = 3, 256, 128, 72   # tests/test_linear_3d.py:100-107 shape
    a = random_csr(m, k, 0.1, seed=8, values="pm1")
    dense = dense_operand((R, k, n), 9, "pm1")
    v, ri, ro, ci = to_dev(a, dev)
    got = ops.left_spmm(m, k, v, ri, ro, ci, torch.from_numpy(dense).to(dev))
    assert tuple(got.shape) == (R, m, n)
    ref = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, dense)
    for r in range(R):
        scale = oracle.spmm_abs(m, k, a.values, a.row_offsets, a.column_indices, dense[r])
        assert_close_fwd(got[r].cpu().numpy(), ref[r], scale, what=f"left replica {r}")
    single = ops.left_spmm(m, k, v, ri, ro, ci, torch.from_numpy(dense[0]).to(dev))
    assert tuple(single.shape) == (1, m, n)


def test_spmm_bias_relu(ops, dev):
    """cuda_spmm.cu.cc:267-277 epilogue; tests/test_spmm_bias_relu.py:27-43 expectation."""
    a, b = spmm_case(72, 64, 72, density=1.0, seed=10, values="u01")
    v, ri, ro, ci = to_dev(a, dev)
    bias = torch.ones(72, device=dev)
    got = ops.spmm_bias(72, 64, v, ri, ro, ci, bias, torch.from_numpy(b).to(dev)).cpu().numpy()
    dense = oracle.dense_spmm(a.to_dense(), b).numpy()
    assert int(((np.abs(got) - np.abs(dense + 1)) < ABS_TORCH).sum()) == 72 * 72
    a, b = spmm_case(130, 90, 200, density=0.2, seed=11, values="pm1")
    bias_np = dense_operand((130,), 12, "pm1")
    v, ri, ro, ci = to_dev(a, dev)
    got = ops.spmm_bias_relu(130, 90, v, ri, ro, ci, torch.from_numpy(bias_np).to(dev),
                             torch.from_numpy(b).to(dev)).cpu().numpy()
    ref = oracle.spmm(130, 90, a.values, a.row_offsets, a.column_indices, b, bias=bias_np, relu=True)
    scale = oracle.spmm_abs(130, 90, a.values, a.row_offsets, a.column_indices, b) + np.abs(bias_np)[:, None]
    assert_close_fwd(got, ref, scale, what="bias+relu")
    assert np.all(got >= 0)


@pytest.mark.parametrize("density", [0.03, 0.15, 0.45, 0.9])
@pytest.mark.parametrize("variant", [1, 2, 3, 4])
def test_spmm_variants_agree(ops, dev, variant, density):
    """Forced kernel variants (1 = gather, 2 = staged, 3/4 = staged with 4/8-row groups) all meet
    the tolerance, including densities where a row overflows the prefetched candidates."""
    a, b = spmm_case(520, 777, 512, density=density, seed=13, values="pm1")
    prev = ops.set_variant("spmm", variant)
    try:
        check_spmm(ops, dev, a, b, f"variant {variant}")
    finally:
        ops.set_variant("spmm", prev)


@pytest.mark.parametrize("density", [0.05, 0.25])
def test_spmm_4096_full_oracle(ops, dev, density):
    """BASELINE.json configs[1] at full size against the fp64 oracle."""
    a, b = spmm_case(4096, 4096, 4096, density=density, seed=0, values="pm1")
    check_spmm(ops, dev, a, b, f"4096^3 d={density}")


@pytest.mark.parametrize("density", [0.1, 0.5])
def test_spmm_4096_properties(ops, dev, density):
    """Size-independent checks at full size: row sums (B = ones), linearity, and a sampled-row oracle."""
    m = k = n = 4096
    a = random_csr(m, k, density, seed=1, values="pm1")
    v, ri, ro, ci = to_dev(a, dev)
    ones = torch.ones((k, n), device=dev)
    got = ops.spmm(m, k, v, ri, ro, ci, ones)
    rows = np.add.reduceat(np.append(a.values.astype(np.float64), 0), a.row_offsets[:-1].clip(max=a.nnz))
    rows[np.diff(a.row_offsets) == 0] = 0
    absrows = np.add.reduceat(np.append(np.abs(a.values).astype(np.float64), 0), a.row_offsets[:-1].clip(max=a.nnz))
    g = got.cpu().numpy()
    assert np.all(g == g[:, :1])                                   # every column identical
    assert np.all(np.abs(g[:, 0] - rows) <= 1e-5 * absrows + 1e-30)
    b1 = torch.from_numpy(dense_operand((k, n), 2, "pm1")).to(dev)
    b2 = torch.from_numpy(dense_operand((k, n), 3, "pm1")).to(dev)
    lin = ops.spmm(m, k, v, ri, ro, ci, b1 + b2) - ops.spmm(m, k, v, ri, ro, ci, b1) - ops.spmm(m, k, v, ri, ro, ci, b2)
    assert float(lin.abs().max()) < 1e-5 * float(absrows.max()) * 4
    # sampled rows against the oracle
    pick = np.random.default_rng(5).choice(m, 64, replace=False)
    sub_ro = np.concatenate([[0], np.cumsum(np.diff(a.row_offsets)[pick])]).astype(np.int32)
    idx = np.concatenate([np.arange(a.row_offsets[i], a.row_offsets[i + 1]) for i in pick])
    b1n = b1.cpu().numpy()
    ref = oracle.spmm(64, k, a.values[idx], sub_ro, a.column_indices[idx], b1n)
    scale = oracle.spmm_abs(64, k, a.values[idx], sub_ro, a.column_indices[idx], b1n)
    assert_close_fwd(ops.spmm(m, k, v, ri, ro, ci, b1)[torch.from_numpy(pick).to(dev)].cpu().numpy(), ref, scale,
                     what="sampled rows")


def test_spmm_through_raw_c_abi(dev):
    """The same kernel through include/tsp.h with plain device pointers (no torch types)."""
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    a, b = spmm_case(200, 300, 136, density=0.1, seed=14, values="pm1")
    v, ri, ro, ci = to_dev(a, dev)
    bd = torch.from_numpy(b).to(dev)
    out = torch.full((200, 136), float("nan"), device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.tsp_spmm_f32(200, 300, 136, a.nnz, 1, 0, p(ri), p(v), p(ro), p(ci), p(bd), None, 0, p(out),
                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, cabi.last_error()
    ref = oracle.spmm(200, 300, a.values, a.row_offsets, a.column_indices, b)
    scale = oracle.spmm_abs(200, 300, a.values, a.row_offsets, a.column_indices, b)
    assert_close_fwd(out.cpu().numpy(), ref, scale, what="c-abi")
    # identity swizzle via NULL row_indices
    out2 = torch.empty_like(out)
    rc = lib.tsp_spmm_f32(200, 300, 136, a.nnz, 1, 0, None, p(v), p(ro), p(ci), p(bd), None, 0, p(out2),
                          ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert torch.equal(out, out2)


def test_spmm_shim_rejects_bad_arguments(ops, dev):
    a, b = spmm_case(16, 16, 16, density=0.5, seed=15)
    v, ri, ro, ci = to_dev(a, dev)
    bd = torch.from_numpy(b).to(dev)
    with pytest.raises(RuntimeError):
        ops.spmm(16, 16, v.double(), ri, ro, ci, bd)
    with pytest.raises(RuntimeError):
        ops.spmm(16, 16, v.cpu(), ri, ro, ci, bd)
    with pytest.raises(RuntimeError):
        ops.spmm(16, 17, v, ri, ro, ci, bd)
    with pytest.raises(RuntimeError):
        ops.spmm(16, 16, v, ri, ro[:-1], ci, bd)


@pytest.mark.parametrize("warps", [4, 16, 32])
@pytest.mark.parametrize("m,k,n,density,reps", [
    (300, 1024, 64, 0.1, 1), (4096, 4096, 64, 0.1, 2), (130, 700, 16, 0.3, 3), (257, 520, 128, 0.2, 1),
    (129, 2048, 100, 0.06, 2), (1000, 1000, 33, 0.15, 1)])
def test_spmm_narrow_operand_wide_ctas(ops, dev, m, k, n, density, reps, warps):
    """Gather variant at the attention P @ V shapes with every CTA width: ragged m / k / n, replicas."""
    a = random_csr(m, k, density, seed=m + k + n, values="pm1")
    lead = (reps,) if reps > 1 else ()
    b = dense_operand(lead + (k, n), 41, "pm1")
    v, ri, ro, ci = to_dev(a, dev)
    vals = torch.from_numpy(np.stack([a.values * (r + 1) for r in range(reps)])).to(dev) if reps > 1 else v
    prev, prevw = ops.set_variant("spmm", 1), ops.set_variant("spmm_warps", warps)
    try:
        got = ops.spmm(m, k, vals, ri, ro, ci, torch.from_numpy(b).to(dev)).cpu().numpy()
    finally:
        ops.set_variant("spmm", prev)
        ops.set_variant("spmm_warps", prevw)
    for r in range(reps):
        vr = a.values * (r + 1) if reps > 1 else a.values
        br = b[r] if reps > 1 else b
        ref = oracle.spmm(m, k, vr, a.row_offsets, a.column_indices, br)
        scale = oracle.spmm_abs(m, k, vr, a.row_offsets, a.column_indices, br)
        assert_close_fwd(got[r] if reps > 1 else got, ref, scale, what=f"narrow {m}x{k}x{n} replica {r}")


@pytest.mark.parametrize("m,k,n,density,reps,variant", [
    (300, 260, 64, 0.1, 3, 0),       # attention backward shape class: gather kernel, 2-D values
    (256, 512, 256, 0.5, 2, 2),      # staged forced: a permuted call still goes through the gather kernel
    (130, 90, 33, 0.3, 1, 0),        # scalar path, 1-D values
])
def test_spmm_perm_multiplies_by_the_transpose(ops, dev, m, k, n, density, reps, variant):
    """spmm_perm(transposed topology, ORIGINAL values, perm) == A^T @ dense: bit-identical to
    spmm on explicitly permuted values, and within the bound of the oracle on the transpose."""
    a = random_csr(m, k, density, seed=5 * m + n, values="pm1")
    lead = (reps,) if reps > 1 else ()
    vals = dense_operand(lead + (a.nnz,), 61, "pm1")
    dense = dense_operand(lead + (m, n), 62, "pm1")
    _, _, ro, ci = to_dev(a, dev)
    tv, td = torch.from_numpy(vals).to(dev), torch.from_numpy(dense).to(dev)
    probe = torch.empty(a.nnz, device=dev)
    _, rot, cit, perm = ops.csr_transpose_with_perm(m, k, probe, ro, ci)
    rit = torch.argsort(rot[1:] - rot[:-1], descending=True, stable=True).to(torch.int32)
    prev = ops.set_variant("spmm", variant)
    try:
        got = ops.spmm_perm(k, m, tv, perm, rit, rot, cit, td)
        explicit = ops.spmm(k, m, ops.permute_values(tv, perm), rit, rot, cit, td)
    finally:
        ops.set_variant("spmm", prev)
    if variant == 0:
        assert torch.equal(got, explicit)       # same (gather) kernel on both sides
    else:
        torch.testing.assert_close(got, explicit, rtol=1e-5, atol=1e-5)
    got = got.cpu().numpy().reshape(lead + (k, n))
    for r in range(reps):
        v_r, d_r = (vals[r], dense[r]) if reps > 1 else (vals, dense)
        vt, rot_o, cit_o, _ = oracle.csr_transpose(m, k, v_r, a.row_offsets, a.column_indices)
        ref = oracle.spmm(k, m, vt, rot_o, cit_o, d_r)
        scale = oracle.spmm_abs(k, m, vt, rot_o, cit_o, d_r)
        assert_close_fwd(got[r] if reps > 1 else got, ref, scale, what=f"spmm_perm replica {r}")
