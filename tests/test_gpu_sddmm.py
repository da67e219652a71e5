"""GPU parity: CUDA SDDMM vs the oracle."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from helpers import ABS_SPUTNIK, assert_close_fwd, to_dev
from oracle.csr import dense_operand, random_csr, row_of

pytestmark = pytest.mark.gpu

# third_party/sputnik/sputnik/sddmm/sddmm_test.cu.cc:76-91  (m, k, n, nnz)
SPUTNIK_SDDMM_PROBLEMS = [
    (1, 32, 32, 32), (1, 64, 32, 32), (1, 32, 64, 64), (1, 32, 64, 32), (1, 32, 64, 16),
    (2, 32, 32, 64), (1024, 80, 1024, 131072), (1, 28, 32, 32), (1024, 128, 1024, 131072),
    (1024, 256, 1024, 131072), (1024, 512, 1024, 131072), (2048, 128, 2048, 1024 * 1024),
    (2048 * 3, 128, 2048, 1024 * 1024 * 3), (2048 * 4, 128, 2048, 1024 * 1024 * 4),
]


def check_sddmm(ops, dev, a, lhs, rhs, what):
    _, ri, ro, ci = to_dev(a, dev)
    got = ops.sddmm(a.m, a.k, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev))
    got = got.cpu().numpy()
    ref = oracle.sddmm(a.m, a.k, a.row_offsets, a.column_indices, lhs, rhs)
    scale = oracle.sddmm_abs(a.m, a.k, a.row_offsets, a.column_indices, lhs, rhs)
    assert got.shape == ref.shape
    assert_close_fwd(got, ref, scale, what=what)
    return got, ref


@pytest.mark.parametrize("m,k,n,nnz", SPUTNIK_SDDMM_PROBLEMS)
def test_sddmm_sputnik_problem_list(ops, dev, m, k, n, nnz):
    a = random_csr(m, n, nnz=nnz, seed=k + n, values="pm1")
    lhs, rhs = dense_operand((m, k), 1, "pm1"), dense_operand((n, k), 2, "pm1")
    got, ref = check_sddmm(ops, dev, a, lhs, rhs, f"sddmm {m}x{k}x{n}/{nnz}")
    assert np.max(np.abs(got - ref)) < ABS_SPUTNIK    # sddmm_test.cu.cc:141


def test_sddmm_reference_test_shape(ops, dev):
    """tests/test_sddmm.py:16-37: 72x64x72, full mask, (lhs @ rhs^T) within 1e-4."""
    a = random_csr(72, 72, 1.0, seed=0, values="u01")
    lhs, rhs = dense_operand((72, 64), 1, "u01"), dense_operand((72, 64), 2, "u01")
    got, _ = check_sddmm(ops, dev, a, lhs, rhs, "72x64x72 full")
    dense = oracle.dense_sddmm(lhs, rhs, np.ones((72, 72), np.float32)).numpy()
    assert np.all(np.abs(got.reshape(72, 72) - dense) < 1e-4 * 4)


@pytest.mark.parametrize("k", [1, 2, 3, 4, 5, 7, 8, 12, 16, 28, 31, 32, 36, 63, 64, 65, 100, 128, 132, 257, 1000])
def test_sddmm_ragged_k(ops, dev, k):
    a = random_csr(45, 77, 0.25, seed=k, values="pm1")
    check_sddmm(ops, dev, a, dense_operand((45, k), 3, "pm1"), dense_operand((77, k), 4, "pm1"), f"k={k}")


@pytest.mark.parametrize("density", [0.0, 0.003, 0.5, 1.0])
def test_sddmm_density_extremes(ops, dev, density):
    a = random_csr(130, 97, density, seed=5, values="pm1")
    if a.nnz == 0:
        _, ri, ro, ci = to_dev(a, dev)
        out = ops.sddmm(130, 97, ri, ro, ci, torch.zeros(130, 64, device=dev), torch.zeros(97, 64, device=dev))
        assert tuple(out.shape) == (0,)
        return
    check_sddmm(ops, dev, a, dense_operand((130, 64), 6, "pm1"), dense_operand((97, 64), 7, "pm1"), f"d={density}")


def test_sddmm_replicated_3d_and_quirks(ops, dev):
    """src/sddmm_cuda.cu:43: [R, nnz] for R > 1, 1-D when the replica count is 1."""
    R, m, n, k = 4, 72, 72, 64
    a = random_csr(m, n, 0.1, seed=8, values="u01", round_to=4)
    lhs, rhs = dense_operand((R, m, k), 9, "pm1"), dense_operand((R, n, k), 10, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    got = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev))
    assert tuple(got.shape) == (R, a.nnz)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    for r in range(R):
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r], rhs[r])
        assert_close_fwd(got[r].cpu().numpy(), ref[r], scale, what=f"replica {r}")
    one = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs[:1]).to(dev), torch.from_numpy(rhs[:1]).to(dev))
    assert tuple(one.shape) == (a.nnz,)


@pytest.mark.parametrize("variant", [1, 2, 3])
def test_sddmm_variants_agree(ops, dev, variant):
    a = random_csr(512, 640, 0.15, seed=11, values="pm1")
    prev = ops.set_variant("sddmm", variant)
    try:
        check_sddmm(ops, dev, a, dense_operand((512, 256), 12, "pm1"), dense_operand((640, 256), 13, "pm1"),
                    f"variant {variant}")
    finally:
        ops.set_variant("sddmm", prev)


@pytest.mark.parametrize("m,n,k,density,reps", [(130, 70, 132, 0.3, 1), (64, 64, 128, 1.0, 1), (300, 513, 260, 0.05, 3),
                                                 (1000, 200, 512, 0.6, 2), (129, 4097, 128, 0.25, 1)])
def test_sddmm_tiled_ragged_shapes(ops, dev, m, n, k, density, reps):
    """Forced tiled variant at shapes that do not divide the 128 x 64 x 32 tile, with replicas."""
    a = random_csr(m, n, density, seed=m + n + k, values="pm1")
    lead = (reps,) if reps > 1 else ()
    lhs, rhs = dense_operand(lead + (m, k), 21, "pm1"), dense_operand(lead + (n, k), 22, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    prev = ops.set_variant("sddmm", 2)
    try:
        got = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)).cpu().numpy()
    finally:
        ops.set_variant("sddmm", prev)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    for r in range(reps):
        g, f = (got[r], ref[r]) if reps > 1 else (got, ref)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r] if reps > 1 else lhs,
                                 rhs[r] if reps > 1 else rhs)
        assert_close_fwd(g, f, scale, what=f"tiled {m}x{n}x{k} replica {r}")


@pytest.mark.parametrize("m,n,k,density,reps,bn", [
    (130, 70, 260, 0.3, 1, 0), (128, 64, 256, 1.0, 1, 0), (300, 513, 260, 0.05, 3, 0), (1000, 200, 512, 0.1, 2, 0),
    (129, 4097, 256, 0.02, 1, 0), (512, 512, 1028, 0.5, 1, 128), (256, 300, 256, 0.9, 1, 256), (640, 384, 300, 0.15, 2, 48)])
def test_sddmm_sampled_ragged_shapes(ops, dev, m, n, k, density, reps, bn):
    """Forced sampled variant: ragged shapes, k tails, replicas, and tiles that overflow the 32*P
    register slots (dense masks / wide forced tiles -> several rounds per tile)."""
    a = random_csr(m, n, density, seed=m + n + k, values="pm1")
    lead = (reps,) if reps > 1 else ()
    lhs, rhs = dense_operand(lead + (m, k), 31, "pm1"), dense_operand(lead + (n, k), 32, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    prev = ops.set_variant("sddmm", 3)
    prev_bn = ops.set_variant("sddmm_bn", bn)
    try:
        out = torch.full(lead + (a.nnz,), float("nan"), device=dev)  # noqa: F841 (poison the allocator's next block)
        del out
        got = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)).cpu().numpy()
    finally:
        ops.set_variant("sddmm", prev)
        ops.set_variant("sddmm_bn", prev_bn)
    assert not np.isnan(got).any()
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    for r in range(reps):
        g, f = (got[r], ref[r]) if reps > 1 else (got, ref)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r] if reps > 1 else lhs,
                                 rhs[r] if reps > 1 else rhs)
        assert_close_fwd(g, f, scale, what=f"sampled {m}x{n}x{k} replica {r}")


def test_sddmm_sampled_empty_rows_and_skew(ops, dev):
    """Rows with no nonzeros, one fully dense row block and an empty column band."""
    m, n, k = 384, 512, 256
    rng = np.random.default_rng(5)
    dense = (rng.random((m, n)) < 0.08)
    dense[10:40] = False
    dense[128:256, :200] = True
    dense[:, 300:360] = False
    from oracle.csr import from_dense_mask
    a = from_dense_mask(dense, seed=6, values="pm1")
    prev = ops.set_variant("sddmm", 3)
    try:
        check_sddmm(ops, dev, a, dense_operand((m, k), 33, "pm1"), dense_operand((n, k), 34, "pm1"), "sampled skew")
    finally:
        ops.set_variant("sddmm", prev)


def test_sddmm_unsorted_columns_fall_back(ops, dev):
    """Shuffled column order inside rows: the tiled kernel must step aside for the gather kernel."""
    a = random_csr(256, 320, 0.4, seed=23, values="pm1")
    rng = np.random.default_rng(3)
    for i in range(a.m):
        s, e = a.row_offsets[i], a.row_offsets[i + 1]
        a.column_indices[s:e] = a.column_indices[s:e][rng.permutation(e - s)]
    check_sddmm(ops, dev, a, dense_operand((256, 256), 24, "pm1"), dense_operand((320, 256), 25, "pm1"), "unsorted")


def test_sddmm_attention_shape(ops, dev):
    """BASELINE.json configs[3] shape per replica: s=4096, d=64, 90 % sparse; 2 replicas here."""
    a = random_csr(4096, 4096, 0.1, seed=14, values="u01")
    lhs, rhs = dense_operand((2, 4096, 64), 15, "pm1"), dense_operand((2, 4096, 64), 16, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    got = ops.sddmm(4096, 4096, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev))
    ref = oracle.sddmm(4096, 4096, a.row_offsets, a.column_indices, lhs, rhs)
    for r in range(2):
        scale = oracle.sddmm_abs(4096, 4096, a.row_offsets, a.column_indices, lhs[r], rhs[r])
        assert_close_fwd(got[r].cpu().numpy(), ref[r], scale, what=f"attention replica {r}")


@pytest.mark.parametrize("density", [0.05, 0.25])
def test_sddmm_4096_full_oracle(ops, dev, density):
    a = random_csr(4096, 4096, density, seed=0, values="pm1")
    check_sddmm(ops, dev, a, dense_operand((4096, 4096), 1, "pm1"), dense_operand((4096, 4096), 2, "pm1"),
                f"4096^3 d={density}")


def test_sddmm_4096_half_density_properties(ops, dev):
    """d = 0.5 at full size: ones x ones = k exactly, and sampled nonzeros against fp64."""
    m = n = k = 4096
    a = random_csr(m, n, 0.5, seed=1, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    out = ops.sddmm(m, n, ri, ro, ci, torch.ones(m, k, device=dev), torch.ones(n, k, device=dev))
    assert bool((out == float(k)).all())
    lhs, rhs = dense_operand((m, k), 3, "pm1"), dense_operand((n, k), 4, "pm1")
    got = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)).cpu().numpy()
    pick = np.random.default_rng(2).choice(a.nnz, 20000, replace=False)
    rows, cols = row_of(a.row_offsets)[pick], a.column_indices[pick]
    l64, r64 = lhs[rows].astype(np.float64), rhs[cols].astype(np.float64)
    ref = np.einsum("ij,ij->i", l64, r64)
    scale = np.einsum("ij,ij->i", np.abs(l64), np.abs(r64))
    assert_close_fwd(got[pick], ref, scale, what="sampled nonzeros")


def test_sddmm_through_raw_c_abi(dev):
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    a = random_csr(150, 170, 0.2, seed=17, values="pm1")
    lhs, rhs = dense_operand((150, 96), 18, "pm1"), dense_operand((170, 96), 19, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    ld, rd = torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)
    out = torch.full((a.nnz,), float("nan"), device=dev)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    rc = lib.tsp_sddmm_f32(150, 96, 170, a.nnz, 1, p(ri), p(ro), p(ci), p(ld), p(rd), p(out),
                           ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, cabi.last_error()
    ref = oracle.sddmm(150, 170, a.row_offsets, a.column_indices, lhs, rhs)
    scale = oracle.sddmm_abs(150, 170, a.row_offsets, a.column_indices, lhs, rhs)
    assert_close_fwd(out.cpu().numpy(), ref, scale, what="c-abi")


@pytest.mark.parametrize("m,n,k,density,reps", [
    (300, 1024, 64, 0.1, 1), (4096, 4096, 64, 0.1, 2), (130, 700, 32, 0.3, 3), (257, 520, 128, 0.2, 1),
    (129, 2048, 96, 0.06, 2), (64, 512, 64, 1.0, 1), (1000, 1000, 64, 0.15, 1)])
@pytest.mark.parametrize("variant", [5])
def test_sddmm_rows_short_inner_dimension(ops, dev, m, n, k, density, reps, variant):
    """Forced rows variant (the attention Q @ K^T shape): ragged m / n, every k in {32..128},
    replicas, dense masks (rows longer than one 32-nonzero batch per band)."""
    a = random_csr(m, n, density, seed=m + n + k, values="pm1")
    lead = (reps,) if reps > 1 else ()
    lhs, rhs = dense_operand(lead + (m, k), 51, "pm1"), dense_operand(lead + (n, k), 52, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    prev = ops.set_variant("sddmm", variant)
    try:
        got = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)).cpu().numpy()
    finally:
        ops.set_variant("sddmm", prev)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    for r in range(reps):
        g, f = (got[r], ref[r]) if reps > 1 else (got, ref)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r] if reps > 1 else lhs,
                                 rhs[r] if reps > 1 else rhs)
        assert_close_fwd(g, f, scale, what=f"rows {m}x{n}x{k} replica {r}")


@pytest.mark.parametrize("variant", [5])
def test_sddmm_rows_empty_rows_and_unsorted(ops, dev, variant):
    m, n, k = 200, 1536, 64
    rng = np.random.default_rng(7)
    mask = rng.random((m, n)) < 0.15
    mask[3:40] = False
    mask[:, 256:768] = False      # two whole bands without a nonzero
    mask[77] = True
    from oracle.csr import from_dense_mask
    a = from_dense_mask(mask, seed=8, values="pm1")
    for i in range(0, m, 3):                      # shuffled columns in every third row: still exact
        s_, e_ = a.row_offsets[i], a.row_offsets[i + 1]
        a.column_indices[s_:e_] = a.column_indices[s_:e_][rng.permutation(e_ - s_)]
    prev = ops.set_variant("sddmm", variant)
    try:
        check_sddmm(ops, dev, a, dense_operand((m, k), 53, "pm1"), dense_operand((n, k), 54, "pm1"), "rows skew")
    finally:
        ops.set_variant("sddmm", prev)


@pytest.mark.parametrize("m,n,k,density,reps", [
    (256, 320, 512, 0.1, 4),      # sampled kernel, replica loop folded into the k loop
    (256, 320, 512, 0.5, 3),      # tiled kernel
    (130, 70, 96, 0.3, 5),        # neither staged kernel applies: scratch + reduce
    (300, 513, 260, 0.05, 2), (128, 128, 256, 0.1, 1),
    (512, 512, 512, 0.1, 4), (512, 2048, 512, 0.1, 4),   # C5 layer shapes: 16 / 64 tiles -> split-K CTAs + partial sums
    (384, 256, 4096, 0.1, 1)])                            # one replica, long k: split-K on the 2-D path
def test_sddmm_reduce_sums_over_replicas(ops, dev, m, n, k, density, reps):
    """sddmm_reduce == sddmm(...).sum(0): the sparse-linear weight gradient (modules/sparse_linear.py:44
    + autograd's broadcast sum) without the [R, nnz] intermediate."""
    a = random_csr(m, n, density, seed=m + n + k + reps, values="pm1")
    lhs, rhs = dense_operand((reps, m, k), 61, "pm1"), dense_operand((reps, n, k), 62, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    got = ops.sddmm_reduce(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev))
    assert tuple(got.shape) == (a.nnz,)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs).astype(np.float64).reshape(reps, -1).sum(0)
    scale = sum(oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r], rhs[r]).astype(np.float64)
                for r in range(reps))
    assert_close_fwd(got.cpu().numpy(), ref, scale, what=f"sddmm_reduce {m}x{n}x{k} R={reps}")


def test_sddmm_reduce_unsorted_columns(ops, dev):
    a = random_csr(256, 320, 0.15, seed=71, values="pm1")
    rng = np.random.default_rng(4)
    for i in range(a.m):
        s, e = a.row_offsets[i], a.row_offsets[i + 1]
        a.column_indices[s:e] = a.column_indices[s:e][rng.permutation(e - s)]
    lhs, rhs = dense_operand((3, 256, 512), 63, "pm1"), dense_operand((3, 320, 512), 64, "pm1")
    _, ri, ro, ci = to_dev(a, dev)
    got = ops.sddmm_reduce(256, 320, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev))
    ref = oracle.sddmm(256, 320, a.row_offsets, a.column_indices, lhs, rhs).astype(np.float64).sum(0)
    scale = sum(oracle.sddmm_abs(256, 320, a.row_offsets, a.column_indices, lhs[r], rhs[r]).astype(np.float64)
                for r in range(3))
    assert_close_fwd(got.cpu().numpy(), ref, scale, what="sddmm_reduce unsorted")


@pytest.mark.parametrize("cfg", [1, 2])
@pytest.mark.parametrize("m,n,k,density,reps", [(1000, 777, 520, 0.3, 1), (256, 192, 132, 0.6, 3), (130, 70, 4096, 0.9, 1)])
def test_sddmm_tiled3_variant_matches_oracle(ops, dev, cfg, m, n, k, density, reps):
    """The k-major dense-tile SDDMM (kernels/sddmm_tiled3.cuh, opt-in through the "sddmm_tiled3" knob): ragged
    shapes, replicas, k not a multiple of the k-step -- against the fp64 oracle within the forward-error bound."""
    a = random_csr(m, n, density, seed=m + k, values="pm1")
    ri, ro, ci = (torch.from_numpy(x).to(dev) for x in (a.row_indices, a.row_offsets, a.column_indices))
    shape_l = (reps, m, k) if reps > 1 else (m, k)
    shape_r = (reps, n, k) if reps > 1 else (n, k)
    lhs, rhs = dense_operand(shape_l, 5, "pm1"), dense_operand(shape_r, 6, "pm1")
    ops.set_variant("sddmm", 2)
    ops.set_variant("sddmm_tiled3", cfg)
    try:
        out = ops.sddmm(m, n, ri, ro, ci, torch.from_numpy(lhs).to(dev), torch.from_numpy(rhs).to(dev)).cpu().numpy()
    finally:
        ops.set_variant("sddmm_tiled3", 0)
        ops.set_variant("sddmm", 0)
    for r in range(reps):
        l, rr = (lhs[r], rhs[r]) if reps > 1 else (lhs, rhs)
        ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, l, rr).astype(np.float64)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, l, rr)
        got = out[r] if reps > 1 else out
        assert np.all(np.abs(got - ref) <= 1e-5 * scale + 1e-30)
