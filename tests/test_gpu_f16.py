"""GPU: 16-bit storage (fp16 / bf16) with fp32 arithmetic -- SURVEY 8 a13 / 8(f) rank 4-iii.

tsp_spmm_f16/bf16, tsp_sddmm_f16/bf16, tsp_csr_softmax_f16/bf16, tsp_permute_values_b16 through the
shim's dtype dispatch, against (a) the fp64 oracle evaluated on the (exactly representable) 16-bit
inputs, (b) tests/golden/ref_b200_f16.npz = outputs of the reference's own half SpMM
(sputnik::CudaSpmm(half2 ...), cuda_spmm.cu.cc:511-570) on a B200, (c) that kernel run live.
Tolerance: one rounding of the RESULT to the storage type (2^-11 relative for fp16, 2^-8 for bf16;
asserted at twice that) + the fp32 accumulation bound 1e-5 * sum|a_i b_i| -- far inside north_star's
1e-2 and Sputnik's own 5e-2 absolute (spmm_test.cu.cc:207)."""
import ctypes
import math
import os

import numpy as np
import pytest
import torch

import oracle
from oracle import ref_gpu
from oracle.csr import dense_operand, even_rows, from_dense_mask, random_csr

pytestmark = pytest.mark.gpu
GOLDEN16 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_b200_f16.npz")
DTYPES = [torch.float16, torch.bfloat16]
EPS = {torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}
TINY = {torch.float16: 2.0 ** -24, torch.bfloat16: 1e-30}     # fp16 results below 6.1e-5 are subnormal


def _t(x, dev):
    return torch.from_numpy(np.ascontiguousarray(x)).to(dev)


def _q(x, dtype, dev):
    """fp32 numpy -> 16-bit tensor on the device, and its exact fp32 image back on the host."""
    t = _t(x, dev).to(dtype)
    return t, t.float().cpu().numpy()


def _check(got, ref64, scale, dtype, what):
    got = got.float().cpu().numpy().astype(np.float64)
    bound = EPS[dtype] * np.abs(ref64) + 1e-5 * np.asarray(scale, dtype=np.float64) + TINY[dtype]
    err = np.abs(got - ref64)
    worst = float(np.max(err / bound)) if err.size else 0.0
    assert worst <= 1.0, f"{what}: max err/bound = {worst:.3f} (max abs err {err.max():.3e})"


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,k,n,density", [(256, 256, 64, 0.1), (72, 64, 72, 0.1), (37, 53, 33, 0.3),
                                           (37, 53, 34, 0.3), (40, 64, 36, 0.3), (128, 96, 512, 0.2),
                                           (64, 64, 8, 0.25), (64, 128, 128, 0.2), (16, 4096, 64, 0.1)])
def test_spmm_16bit_matches_oracle(ops, dev, dtype, m, k, n, density):
    a = random_csr(m, k, density, seed=m + n, values="pm1")
    vals, vals32 = _q(a.values, dtype, dev)
    dense, dense32 = _q(dense_operand((k, n), 3, "pm1"), dtype, dev)
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    out = ops.spmm(m, k, vals, ri, ro, ci, dense)
    assert out.dtype == dtype and tuple(out.shape) == (m, n)
    ref = oracle.spmm(m, k, vals32, a.row_offsets, a.column_indices, dense32).astype(np.float64)
    scale = oracle.spmm_abs(m, k, vals32, a.row_offsets, a.column_indices, dense32)
    _check(out, ref, scale, dtype, f"spmm {dtype}")


@pytest.mark.parametrize("dtype", DTYPES)
def test_spmm_16bit_replicas_bias_relu_perm_and_empty_rows(ops, dev, dtype):
    m, k, n, reps = 96, 80, 64, 3
    a = random_csr(m, k, 0.15, seed=7, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    v2, v2_32 = _q(dense_operand((reps, a.nnz), 8, "pm1"), dtype, dev)
    d3, d3_32 = _q(dense_operand((reps, k, n), 9, "pm1"), dtype, dev)
    out = ops.spmm(m, k, v2, ri, ro, ci, d3)                       # values [R, nnz]
    assert tuple(out.shape) == (reps, m, n) and out.dtype == dtype
    for r in range(reps):
        ref = oracle.spmm(m, k, v2_32[r], a.row_offsets, a.column_indices, d3_32[r]).astype(np.float64)
        scale = oracle.spmm_abs(m, k, v2_32[r], a.row_offsets, a.column_indices, d3_32[r])
        _check(out[r], ref, scale, dtype, "spmm 3-D")
    left = ops.left_spmm(m, k, v2[0].contiguous(), ri, ro, ci, d3)  # one values vector for every replica
    for r in range(reps):
        ref = oracle.spmm(m, k, v2_32[0], a.row_offsets, a.column_indices, d3_32[r]).astype(np.float64)
        scale = oracle.spmm_abs(m, k, v2_32[0], a.row_offsets, a.column_indices, d3_32[r])
        _check(left[r], ref, scale, dtype, "left_spmm")
    bias = dense_operand((m,), 10, "pm1")
    got = ops.spmm_bias_relu(m, k, v2[0].contiguous(), ri, ro, ci, _t(bias, dev), d3[0].contiguous())
    ref = oracle.spmm(m, k, v2_32[0], a.row_offsets, a.column_indices, d3_32[0], bias=bias, relu=True).astype(np.float64)
    scale = oracle.spmm_abs(m, k, v2_32[0], a.row_offsets, a.column_indices, d3_32[0]) + np.abs(bias)[:, None]
    _check(got, ref, scale, dtype, "spmm_bias_relu")
    # A^T @ G through the transposed topology and the permutation (what the backward passes use)
    rot, cit, perm, rit = ops.transposed_topology(m, k, ro, ci)
    gmat, g32 = _q(dense_operand((m, n), 11, "pm1"), dtype, dev)
    got = ops.spmm_perm(k, m, v2[0].contiguous(), perm, rit, rot, cit, gmat)
    dmat = np.zeros((m, k))
    dmat[np.repeat(np.arange(m), np.diff(a.row_offsets)), a.column_indices] = v2_32[0]
    _check(got, dmat.T @ g32.astype(np.float64), np.abs(dmat).T @ np.abs(g32).astype(np.float64), dtype, "spmm_perm")
    # csr_transpose moves 16-bit values exactly (planned and unplanned)
    _, _, _, eperm = oracle.csr_transpose(m, k, a.values, a.row_offsets, a.column_indices)
    for policy in (1, 0):
        old = ops.plan_policy(policy)
        vt, rot2, cit2 = ops.csr_transpose(m, k, v2, ro, ci)
        ops.plan_policy(old)
        assert vt.dtype == dtype and torch.equal(vt, v2[:, _t(eperm, dev).long()])
        assert torch.equal(rot2, rot) and torch.equal(cit2, cit)
    # rows without nonzeros produce zeros
    mask = np.zeros((m, k), dtype=bool)
    mask[np.repeat(np.arange(m), np.diff(a.row_offsets)), a.column_indices] = True
    mask[2::3] = False
    e = from_dense_mask(mask, seed=12, values="pm1")
    ve, ve32 = _q(e.values, dtype, dev)
    out = ops.spmm(m, k, ve, _t(e.row_indices, dev), _t(e.row_offsets, dev), _t(e.column_indices, dev), d3[0].contiguous())
    ref = oracle.spmm(m, k, ve32, e.row_offsets, e.column_indices, d3_32[0]).astype(np.float64)
    _check(out, ref, oracle.spmm_abs(m, k, ve32, e.row_offsets, e.column_indices, d3_32[0]), dtype, "spmm, empty rows")
    assert float(out[2].abs().max()) == 0.0


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,n,k,density", [(256, 256, 64, 0.1), (128, 96, 128, 0.2), (64, 80, 192, 0.2),
                                           (64, 64, 256, 0.3), (72, 72, 72, 0.3), (40, 48, 33, 0.3),
                                           (32, 32, 4096, 0.5), (37, 53, 16, 0.4)])
def test_sddmm_16bit_matches_oracle(ops, dev, dtype, m, n, k, density):
    a = random_csr(m, n, density, seed=k, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    lhs, lhs32 = _q(dense_operand((m, k), 5, "pm1"), dtype, dev)
    rhs, rhs32 = _q(dense_operand((n, k), 6, "pm1"), dtype, dev)
    out = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
    assert out.dtype == dtype and tuple(out.shape) == (a.nnz,)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs32, rhs32).astype(np.float64)
    scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs32, rhs32)
    _check(out, ref, scale, dtype, f"sddmm {dtype} k={k}")
    # replicated: [R, nnz]
    l3, l3_32 = _q(dense_operand((2, m, k), 7, "pm1"), dtype, dev)
    r3, r3_32 = _q(dense_operand((2, n, k), 8, "pm1"), dtype, dev)
    out = ops.sddmm(m, n, ri, ro, ci, l3, r3)
    assert tuple(out.shape) == (2, a.nnz)
    for r in range(2):
        ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, l3_32[r], r3_32[r]).astype(np.float64)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, l3_32[r], r3_32[r])
        _check(out[r], ref, scale, dtype, "sddmm 3-D")


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,n,density,reps", [(72, 72, 0.1, 1), (16, 4096, 0.5, 2), (128, 4096, 0.1, 3)])
def test_softmax_16bit_matches_oracle(ops, dev, dtype, m, n, density, reps):
    a = random_csr(m, n, density, seed=n + reps, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    x, x32 = _q(dense_operand((reps, a.nnz), 12, "pm1") * 6.0, dtype, dev)
    for scale in (1.0, 0.125):
        got = ops.sparse_softmax_scaled(x, ri, ro, ci, scale) if scale != 1.0 else ops.sparse_softmax(x, ri, ro, ci)
        assert got.dtype == dtype and got.shape == x.shape
        for r in range(reps):
            ref = oracle.csr_softmax(x32[r], a.row_offsets, scale=scale).astype(np.float64)
            _check(got[r], ref, 0.0 * ref, dtype, f"softmax {dtype}")


@pytest.mark.parametrize("dtype", DTYPES)
def test_attention_chain_16bit_within_north_star_tolerance(ops, dev, dtype):
    """sddmm -> softmax(scale) -> spmm in 16-bit storage vs dense fp64 masked attention: 1e-2 (north_star)."""
    reps, s, d = 4, 512, 64
    a = random_csr(s, s, 0.1, seed=21, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    q, k, v = (torch.randn(reps, s, d, device=dev).to(dtype) for _ in range(3))
    scores = ops.sddmm(s, s, ri, ro, ci, q, k)
    probs = ops.sparse_softmax_scaled(scores, ri, ro, ci, 1.0 / math.sqrt(d))
    out = ops.spmm(s, s, probs, ri, ro, ci, v)
    mask = torch.zeros(s, s, dtype=torch.bool, device=dev)
    mask[torch.repeat_interleave(torch.arange(s, device=dev), (ro[1:] - ro[:-1]).long()), ci.long()] = True
    sc = (q.double() @ k.double().transpose(-2, -1)) / math.sqrt(d)
    expect = torch.softmax(sc.masked_fill(~mask, float("-inf")), dim=-1) @ v.double()
    torch.testing.assert_close(out.double(), expect, rtol=1e-2, atol=1e-2)


needs_golden = pytest.mark.skipif(not os.path.exists(GOLDEN16), reason="fp16 golden vectors not generated yet")
needs_ref = pytest.mark.skipif(not ref_gpu.available(), reason="oracle/_ref/libsputnik_ref.so not built")


def _cases(g):
    return sorted({k.split("_")[1] for k in g.files if k.startswith("spmmh_")})


@needs_golden
def test_spmm_f16_matches_the_reference_half_kernel_golden(ops, dev):
    g = np.load(GOLDEN16)
    assert len(_cases(g)) >= 8
    for c in _cases(g):
        p = f"spmmh_{c}_"
        m, k, n = (int(x) for x in g[p + "shape"])
        args = (m, k, _t(g[p + "values"], dev), _t(g[p + "row_indices"], dev), _t(g[p + "row_offsets"], dev),
                _t(g[p + "column_indices"], dev))
        bias = g[p + "bias"] if p + "bias" in g.files else None
        got = ops.spmm_bias_relu(*args, _t(bias, dev), _t(g[p + "dense"], dev)) if bias is not None else \
            ops.spmm(*args, _t(g[p + "dense"], dev))
        v32, d32 = g[p + "values"].astype(np.float32), g[p + "dense"].astype(np.float32)
        scale = oracle.spmm_abs(m, k, v32, g[p + "row_offsets"], g[p + "column_indices"], d32)
        if bias is not None:
            scale = scale + np.abs(bias)[:, None]
        gold = g[p + "out"].astype(np.float64)
        # both sides round an fp32-accumulated sum to fp16: at most one fp16 ulp apart + accumulation-order noise
        _check(got, gold, 2.0 * scale, torch.float16, f"vs sputnik half golden {c}")


@needs_ref
def test_spmm_f16_matches_the_reference_half_kernel_live(ops, dev):
    for seed, (m, k, n, d) in enumerate([(512, 512, 64, 0.1), (1024, 1024, 128, 0.125), (256, 2048, 256, 0.05)]):
        a = even_rows(random_csr(m, k, d, seed=40 + seed, values="pm1"))
        vals, v32 = _q(a.values, torch.float16, dev)
        dense, d32 = _q(dense_operand((k, n), 50 + seed, "pm1"), torch.float16, dev)
        ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
        theirs = ref_gpu.spmm_half(m, k, vals, ri, ro, ci, dense)
        ours = ops.spmm(m, k, vals, ri, ro, ci, dense)
        scale = oracle.spmm_abs(m, k, v32, a.row_offsets, a.column_indices, d32)
        _check(ours, theirs.float().cpu().numpy().astype(np.float64), 2.0 * scale, torch.float16, "vs sputnik half live")
        ref = oracle.spmm(m, k, v32, a.row_offsets, a.column_indices, d32).astype(np.float64)
        _check(ours, ref, scale, torch.float16, "vs oracle")


def test_16bit_entry_points_through_the_raw_c_abi(ops, dev):
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    m, k, n = 128, 96, 64
    a = random_csr(m, k, 0.2, seed=60, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    for dtype, spmm_fn, sddmm_fn, sm_fn in ((torch.float16, lib.tsp_spmm_f16, lib.tsp_sddmm_f16, lib.tsp_csr_softmax_f16),
                                            (torch.bfloat16, lib.tsp_spmm_bf16, lib.tsp_sddmm_bf16, lib.tsp_csr_softmax_bf16)):
        vals, _ = _q(a.values, dtype, dev)
        dense, _ = _q(dense_operand((k, n), 61, "pm1"), dtype, dev)
        out = torch.empty(m, n, dtype=dtype, device=dev)
        rc = spmm_fn(m, k, n, a.nnz, 1, 0, ri.data_ptr(), vals.data_ptr(), None, ro.data_ptr(), ci.data_ptr(),
                     dense.data_ptr(), None, 0, out.data_ptr(), stream)
        assert rc == 0, cabi.last_error()
        assert torch.equal(out, ops.spmm(m, k, vals, ri, ro, ci, dense))
        lhs, _ = _q(dense_operand((m, 64), 62, "pm1"), dtype, dev)
        rhs, _ = _q(dense_operand((k, 64), 63, "pm1"), dtype, dev)
        sc = torch.empty(a.nnz, dtype=dtype, device=dev)
        rc = sddmm_fn(m, 64, k, a.nnz, 1, None, ro.data_ptr(), ci.data_ptr(), lhs.data_ptr(), rhs.data_ptr(),
                      sc.data_ptr(), stream)
        assert rc == 0, cabi.last_error()
        assert torch.equal(sc, ops.sddmm(m, k, ri, ro, ci, lhs, rhs))
        pr = torch.empty_like(sc)
        rc = sm_fn(m, a.nnz, 1, ctypes.c_float(0.5), sc.data_ptr(), None, ro.data_ptr(), pr.data_ptr(), stream)
        assert rc == 0, cabi.last_error()
        assert torch.equal(pr, ops.sparse_softmax_scaled(sc, ri, ro, ci, 0.5))
    assert spmm_fn(-1, 1, 1, 0, 1, 0, None, None, None, None, None, None, None, 0, None, None) == 1


def test_mixed_dtypes_are_rejected(ops, dev):
    a = random_csr(16, 16, 0.5, seed=1, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    vals = _t(a.values, dev)
    with pytest.raises(RuntimeError, match="same dtype"):
        ops.spmm(16, 16, vals.half(), ri, ro, ci, torch.zeros(16, 8, device=dev))
    with pytest.raises(RuntimeError, match="same dtype"):
        ops.sddmm(16, 16, ri, ro, ci, torch.zeros(16, 8, device=dev).half(), torch.zeros(16, 8, device=dev).bfloat16())
    with pytest.raises(RuntimeError, match="float32 only"):
        ops.sddmm_reduce(16, 16, ri, ro, ci, torch.zeros(2, 16, 8, device=dev).half(), torch.zeros(2, 16, 8, device=dev).half())


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("density", [0.4, 0.08])
def test_16bit_storage_through_the_fp32_staged_kernels(ops, dev, dtype, density):
    """Big, dense-ish shapes: the 16-bit operands are widened once, the fp32 staged kernel (merged-row SpMM, tiled /
    sampled SDDMM) runs and the result is rounded once -- same contract, checked against the fp64 oracle; and the
    result equals the fp32 op on the widened operands rounded to the storage type, bit for bit."""
    m, k, n = 1024, 1024, 512
    a = random_csr(m, k, density, seed=70, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    vals, v32 = _q(a.values, dtype, dev)
    dense, d32 = _q(dense_operand((k, n), 71, "pm1"), dtype, dev)
    out = ops.spmm(m, k, vals, ri, ro, ci, dense)
    assert out.dtype == dtype
    ref = oracle.spmm(m, k, v32, a.row_offsets, a.column_indices, d32).astype(np.float64)
    _check(out, ref, oracle.spmm_abs(m, k, v32, a.row_offsets, a.column_indices, d32), dtype, "spmm, widened")
    if density >= 0.35:
        assert torch.equal(out, ops.spmm(m, k, vals.float(), ri, ro, ci, dense.float()).to(dtype))
    lhs, l32 = _q(dense_operand((m, n), 72, "pm1"), dtype, dev)
    rhs, r32 = _q(dense_operand((k, n), 73, "pm1"), dtype, dev)
    sc = ops.sddmm(m, k, ri, ro, ci, lhs, rhs)
    assert sc.dtype == dtype
    sref = oracle.sddmm(m, k, a.row_offsets, a.column_indices, l32, r32).astype(np.float64)
    _check(sc, sref, oracle.sddmm_abs(m, k, a.row_offsets, a.column_indices, l32, r32), dtype, "sddmm, widened")
    ops.set_variant("sddmm_b16", 1)   # (never the native 16-bit sampled kernel: its sums associate differently)
    try:
        assert torch.equal(ops.sddmm(m, k, ri, ro, ci, lhs, rhs), ops.sddmm(m, k, ri, ro, ci, lhs.float(), rhs.float()).to(dtype))
    finally:
        ops.set_variant("sddmm_b16", 0)


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,n,k,density,reps,sort", [(2048, 2048, 1024, 0.1, 1, True), (1024, 1536, 520, 0.05, 3, True),
                                                     (2048, 2048, 512, 0.2, 1, False)])
def test_sddmm_16bit_native_sampled_kernel(ops, dev, dtype, m, n, k, density, reps, sort):
    """Masks of 1.5 % .. 40 % with a long inner dimension: the TMA-staged sampled kernel on the 16-bit operands
    themselves (packed slabs, mixed-precision FMA, fp32 accumulators, one rounding) -- against the fp64 oracle, against
    the widened fp32 path, with a k that is not a multiple of the 64-element slab, and with unsorted columns (the
    device guard hands the work to the gather kernel)."""
    a = random_csr(m, n, density, seed=k + m, values="pm1")
    cols = a.column_indices.copy()
    if not sort:   # reverse the columns of every row: same mask, descending order
        for r in range(m):
            cols[a.row_offsets[r]:a.row_offsets[r + 1]] = cols[a.row_offsets[r]:a.row_offsets[r + 1]][::-1]
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(cols, dev)
    shape_l = (reps, m, k) if reps > 1 else (m, k)
    shape_r = (reps, n, k) if reps > 1 else (n, k)
    lhs, l32 = _q(dense_operand(shape_l, 5, "pm1"), dtype, dev)
    rhs, r32 = _q(dense_operand(shape_r, 6, "pm1"), dtype, dev)
    out = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)   # plans the topology: sortedness check + block-start view
    assert out.dtype == dtype
    torch.cuda.synchronize()
    before = ops.launch_count()
    again = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
    launches = ops.launch_count() - before
    assert torch.equal(again, out)
    ops.set_variant("sddmm_b16", 1)
    try:
        wide = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
    finally:
        ops.set_variant("sddmm_b16", 0)
    for r in range(reps):
        lr, rr = (l32[r], r32[r]) if reps > 1 else (l32, r32)
        ref = oracle.sddmm(m, n, a.row_offsets, cols, lr, rr).astype(np.float64)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, cols, lr, rr)
        _check(out[r] if reps > 1 else out, ref, scale, dtype, f"native sampled sddmm {dtype} replica {r}")
        _check(wide[r] if reps > 1 else wide, ref, scale, dtype, f"widened sddmm {dtype} replica {r}")
    if sort:
        assert launches == 1, launches   # verdict known, view cached: the sampled kernel alone
