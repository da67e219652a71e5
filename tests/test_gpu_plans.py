"""GPU: per-topology plans (tsp_plan_*; the shim's plan cache) -- planned and unplanned calls agree
bit for bit, the cache cannot serve a stale topology, a planned call launches exactly one kernel."""
import ctypes

import numpy as np
import pytest
import torch

import oracle
from helpers import assert_close_fwd, spmm_case, to_dev
from oracle.csr import dense_operand, random_csr

pytestmark = pytest.mark.gpu


@pytest.fixture(autouse=True)
def _fresh_cache(ops):
    ops.plan_cache_clear()
    old = ops.plan_policy(1)
    yield
    ops.plan_policy(old)
    ops.set_variant("spmm", 0)
    ops.set_variant("sddmm", 0)


def _t(x, dev):
    return torch.from_numpy(np.ascontiguousarray(x)).to(dev)


@pytest.mark.parametrize("m,k,n,density", [(1024, 1024, 256, 0.3), (512, 768, 256, 0.08), (96, 80, 72, 0.2)])
def test_planned_spmm_equals_unplanned(ops, dev, m, k, n, density):
    a, b = spmm_case(m, k, n, density, seed=3)
    v, ri, ro, ci = to_dev(a, dev)
    bd = _t(b, dev)
    ops.plan_policy(0)
    base = ops.spmm(m, k, v, ri, ro, ci, bd)
    ops.plan_policy(1)
    _, hits0, misses0, _ = ops.plan_cache_stats()
    first = ops.spmm(m, k, v, ri, ro, ci, bd)          # plan created, verdict maybe not on the host yet
    torch.cuda.synchronize()
    before = ops.launch_count()
    second = ops.spmm(m, k, v, ri, ro, ci, bd)         # verdict known: the operator's kernel and nothing else
    assert ops.launch_count() - before == 1
    assert torch.equal(base, first) and torch.equal(base, second)
    ref = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, b)
    scale = oracle.spmm_abs(m, k, a.values, a.row_offsets, a.column_indices, b)
    assert_close_fwd(second.cpu().numpy(), ref, scale, what="planned spmm")
    entries, hits, misses, _ = ops.plan_cache_stats()
    assert entries == 1 and hits - hits0 == 1 and misses - misses0 == 1


# (the sampled kernel splits K over several CTAs per tile on masks this small: + one sum launch)
@pytest.mark.parametrize("m,n,k,density,variant,launches", [(512, 512, 512, 0.3, 0, 1), (1024, 1024, 512, 0.08, 0, 2),
                                                            (512, 512, 64, 0.1, 0, 1), (1024, 1024, 256, 0.1, 3, 2),
                                                            (1024, 1024, 256, 0.3, 2, 1)])
def test_planned_sddmm_equals_unplanned(ops, dev, m, n, k, density, variant, launches):
    a = random_csr(m, n, density, seed=4, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    lhs, rhs = dense_operand((m, k), 5, "pm1"), dense_operand((n, k), 6, "pm1")
    ld, rd = _t(lhs, dev), _t(rhs, dev)
    ops.set_variant("sddmm", variant)
    ops.plan_policy(0)
    base = ops.sddmm(m, n, ri, ro, ci, ld, rd)
    ops.plan_policy(1)
    first = ops.sddmm(m, n, ri, ro, ci, ld, rd)
    torch.cuda.synchronize()
    before = ops.launch_count()
    second = ops.sddmm(m, n, ri, ro, ci, ld, rd)
    assert ops.launch_count() - before <= launches
    assert torch.equal(base, first) and torch.equal(base, second)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs, rhs)
    assert_close_fwd(second.cpu().numpy(), ref, scale, what="planned sddmm")


def test_planned_sddmm_reduce(ops, dev):
    m, n, k, reps = 512, 512, 256, 8
    a = random_csr(m, n, 0.1, seed=7, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    lhs, rhs = dense_operand((reps, m, k), 8, "pm1"), dense_operand((reps, n, k), 9, "pm1")
    ld, rd = _t(lhs, dev), _t(rhs, dev)
    ops.plan_policy(0)
    base = ops.sddmm_reduce(m, n, ri, ro, ci, ld, rd)
    ops.plan_policy(1)
    ops.sddmm_reduce(m, n, ri, ro, ci, ld, rd)
    torch.cuda.synchronize()
    got = ops.sddmm_reduce(m, n, ri, ro, ci, ld, rd)
    assert torch.equal(base, got)
    ref = sum(oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs[r], rhs[r]).astype(np.float64) for r in range(reps))
    scale = sum(oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs[r], rhs[r]) for r in range(reps))
    assert_close_fwd(got.cpu().numpy(), ref, scale, what="planned sddmm_reduce")


def test_planned_unsorted_columns_take_the_gather_kernels(ops, dev):
    """Columns descending inside every row: the plan's verdict routes both ops to the gather kernels."""
    m, k, n = 1024, 1024, 256
    a, b = spmm_case(m, k, n, 0.3, seed=10)
    cols, vals = a.column_indices.copy(), a.values.copy()
    for r in range(m):
        s, e = a.row_offsets[r], a.row_offsets[r + 1]
        cols[s:e] = cols[s:e][::-1]
        vals[s:e] = vals[s:e][::-1]
    ri, ro = _t(a.row_indices, dev), _t(a.row_offsets, dev)
    ci, v, bd = _t(cols, dev), _t(vals, dev), _t(b, dev)
    ref = oracle.spmm(m, k, vals, a.row_offsets, cols, b)
    scale = oracle.spmm_abs(m, k, vals, a.row_offsets, cols, b)
    for _ in range(2):
        got = ops.spmm(m, k, v, ri, ro, ci, bd)
        torch.cuda.synchronize()
        assert_close_fwd(got.cpu().numpy(), ref, scale, what="spmm, unsorted, planned")
    lhs, rhs = dense_operand((m, 512), 11, "pm1"), dense_operand((k, 512), 12, "pm1")
    sref = oracle.sddmm(m, k, a.row_offsets, cols, lhs, rhs)
    sscale = oracle.sddmm_abs(m, k, a.row_offsets, cols, lhs, rhs)
    for _ in range(2):
        got = ops.sddmm(m, k, ri, ro, ci, _t(lhs, dev), _t(rhs, dev))
        torch.cuda.synchronize()
        assert_close_fwd(got.cpu().numpy(), sref, sscale, what="sddmm, unsorted, planned")


def test_plan_cache_is_not_fooled_by_a_recycled_address(ops, dev):
    """same row_offsets, a NEW column_indices of the same length every step (top-k / random
    block masks).  The caching allocator would hand the freed address back; the cache holds the tensor
    it planned, so the address cannot be recycled, and results follow the new topology."""
    m, k, n, per_row = 256, 256, 128, 32
    ro = torch.arange(0, m * per_row + 1, per_row, dtype=torch.int32, device=dev)
    ri = torch.arange(m, dtype=torch.int32, device=dev)
    g = torch.Generator(device="cpu").manual_seed(0)
    dense = torch.randn(k, n, device=dev)
    vals = torch.randn(m * per_row, device=dev)
    for step in range(6):
        cols = torch.stack([torch.randperm(k, generator=g)[:per_row].sort().values for _ in range(m)])
        ci = cols.reshape(-1).to(torch.int32).to(dev)           # fresh tensor, same size, every step
        out = ops.spmm(m, k, vals, ri, ro, ci, dense)
        dmat = torch.zeros(m, k, device=dev, dtype=torch.float64)
        dmat[torch.arange(m, device=dev).repeat_interleave(per_row), ci.long()] = vals.double()
        torch.testing.assert_close(out.double(), dmat @ dense.double(), rtol=1e-5, atol=1e-5)
        t = ops.transposed_topology(m, k, ro, ci)
        dt = (dmat != 0).t().contiguous().float().to_sparse_csr()
        assert torch.equal(t[0].long(), dt.crow_indices()) and torch.equal(t[1].long(), dt.col_indices())
        del ci, out, t


def test_plan_cache_sees_in_place_edits(ops, dev):
    m, k, n = 128, 128, 64
    a, b = spmm_case(m, k, n, 0.2, seed=13)
    v, ri, ro, ci = to_dev(a, dev)
    bd = _t(b, dev)
    ops.spmm(m, k, v, ri, ro, ci, bd)
    t0 = [x.clone() for x in ops.transposed_topology(m, k, ro, ci)]
    # shift every column by one (mod k), keeping rows sorted where possible: a different topology, same tensors
    new_cols = ((ci.long() + 1) % k).to(torch.int32)
    ci.copy_(new_cols)                                           # in place: _version changes
    out = ops.spmm(m, k, v, ri, ro, ci, bd)
    dmat = torch.zeros(m, k, device=dev, dtype=torch.float64)
    rows = torch.repeat_interleave(torch.arange(m, device=dev), (ro[1:] - ro[:-1]).long())
    dmat[rows, ci.long()] = v.double()
    torch.testing.assert_close(out.double(), dmat @ bd.double(), rtol=1e-5, atol=1e-5)
    t1 = ops.transposed_topology(m, k, ro, ci)
    assert not torch.equal(t0[1], t1[1])


def test_csr_transpose_planned_is_bit_exact_and_cached(ops, dev):
    m, n = 300, 200
    a = random_csr(m, n, 0.15, seed=14, values="pm1")
    v, _, ro, ci = to_dev(a, dev)
    evt, erot, ecit, eperm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    ops.plan_policy(0)
    u = ops.csr_transpose(m, n, v, ro, ci)
    ops.plan_policy(1)
    p1 = ops.csr_transpose(m, n, v, ro, ci)
    before = ops.launch_count()
    p2 = ops.csr_transpose_with_perm(m, n, v, ro, ci)
    assert ops.launch_count() - before == 1                      # one gather through the cached permutation
    for got in (u, p1, p2):
        assert np.array_equal(got[0].cpu().numpy(), evt)
        assert np.array_equal(got[1].cpu().numpy(), erot) and np.array_equal(got[2].cpu().numpy(), ecit)
    assert np.array_equal(p2[3].cpu().numpy(), eperm)
    assert p1[1].data_ptr() == p2[1].data_ptr()                  # the plan's own arrays, shared between calls
    vr = _t(dense_operand((3, a.nnz), 15, "pm1"), dev)           # replicated values through the same plan
    got = ops.csr_transpose(m, n, vr, ro, ci)[0]
    assert torch.equal(got, vr[:, _t(eperm, dev).long()])
    with pytest.raises(RuntimeError, match="modified in place"):
        p1[1].add_(1)
        ops.csr_transpose(m, n, v, ro, ci)
    ops.plan_cache_clear()


def test_plan_through_the_raw_c_abi(ops, dev):
    from torch_sputnik_b200 import cabi
    lib = cabi.load()
    m, k, n = 384, 512, 256
    a, b = spmm_case(m, k, n, 0.3, seed=16)
    v, ri, ro, ci = to_dev(a, dev)
    bd = _t(b, dev)
    stream = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    plan = ctypes.c_void_p()
    assert lib.tsp_plan_create(m, k, a.nnz, ro.data_ptr(), ci.data_ptr(), stream, ctypes.byref(plan)) == 0
    torch.cuda.synchronize()
    assert lib.tsp_plan_sorted(plan) == 1
    out = torch.empty(m, n, device=dev)
    rc = lib.tsp_plan_spmm_f32(plan, n, 1, 0, ri.data_ptr(), v.data_ptr(), None, bd.data_ptr(), None, 0,
                               out.data_ptr(), stream)
    assert rc == 0, cabi.last_error()
    ops.plan_policy(0)
    assert torch.equal(out, ops.spmm(m, k, v, ri, ro, ci, bd))
    # A^T @ G through the transposed child plan and the permutation, no values_t materialised
    child, ro_t, ci_t, perm = (ctypes.c_void_p() for _ in range(4))
    rc = lib.tsp_plan_transpose(plan, stream, ctypes.byref(child), ctypes.byref(ro_t), ctypes.byref(ci_t),
                                ctypes.byref(perm))
    assert rc == 0, cabi.last_error()
    gmat = _t(dense_operand((m, n), 17, "pm1"), dev)
    outt = torch.empty(k, n, device=dev)
    rc = lib.tsp_plan_spmm_f32(child, n, 1, 0, None, v.data_ptr(), perm, gmat.data_ptr(), None, 0,
                               outt.data_ptr(), stream)
    assert rc == 0, cabi.last_error()
    dmat = torch.zeros(m, k, device=dev, dtype=torch.float64)
    rows = torch.repeat_interleave(torch.arange(m, device=dev), (ro[1:] - ro[:-1]).long())
    dmat[rows, ci.long()] = v.double()
    torch.testing.assert_close(outt.double(), dmat.t() @ gmat.double(), rtol=1e-5, atol=1e-5)
    # planned sddmm
    lhs, rhs = _t(dense_operand((m, 256), 18, "pm1"), dev), _t(dense_operand((k, 256), 19, "pm1"), dev)
    sout = torch.empty(a.nnz, device=dev)
    rc = lib.tsp_plan_sddmm_f32(plan, 256, 1, 0, None, lhs.data_ptr(), rhs.data_ptr(), sout.data_ptr(), stream)
    assert rc == 0, cabi.last_error()
    assert torch.equal(sout, ops.sddmm(m, k, ri, ro, ci, lhs, rhs))
    torch.cuda.synchronize()
    lib.tsp_plan_destroy(plan)


@pytest.mark.parametrize("variant", [1, 3, 4])
def test_spmm_non_finite_operand_reaches_only_the_rows_that_own_the_column(ops, dev, variant):
    """the merged-row staged SpMM must not turn an Inf/NaN in dense row c into NaN in rows
    that merely share a group with a row holding column c (cuda_spmm.cu.cc:224-238 touches only the
    dense rows a row names).  Variants: 1 gather, 3 staged 4-row groups, 4 staged 8-row groups."""
    m, k, n = 512, 512, 384
    a, b = spmm_case(m, k, n, 0.3, seed=20)
    b = b.copy()
    b[7, 5] = np.inf
    b[100, :] = np.nan
    b[301, 200] = -np.inf
    v, ri, ro, ci = to_dev(a, dev)
    ops.set_variant("spmm", variant)
    out = ops.spmm(m, k, v, ri, ro, ci, _t(b, dev)).cpu().numpy()
    dense_a = np.zeros((m, k), dtype=np.float64)
    rows = np.repeat(np.arange(m), np.diff(a.row_offsets))
    dense_a[rows, a.column_indices] = a.values
    with np.errstate(invalid="ignore", over="ignore"):
        bad = np.zeros((m, n), dtype=bool)
        for c in (7, 100, 301):
            bad |= np.outer(dense_a[:, c] != 0, ~np.isfinite(b[c]))
    assert np.array_equal(~np.isfinite(out), bad)
    b_fin = np.where(np.isfinite(b), b, 0.0).astype(np.float32)
    ref = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, b_fin)
    scale = oracle.spmm_abs(m, k, a.values, a.row_offsets, a.column_indices, b_fin)
    good = ~bad
    assert_close_fwd(out[good], ref[good], scale[good], what=f"spmm variant {variant}, finite part")


def test_planned_ops_inside_a_cuda_graph(ops, dev):
    """(a) a topology planned just before the capture, its verdict not yet on the host: no event is queried
    under capture, the call takes the device-guarded path; (b) a topology first seen DURING the capture: no
    plan is created (cudaMalloc is illegal there), the unplanned entry runs.  Replays give the eager result."""
    m, k, n = 1024, 1024, 256
    a, b = spmm_case(m, k, n, 0.3, seed=21)
    a2, b2 = spmm_case(m, k, n, 0.08, seed=22)
    v, ri, ro, ci = to_dev(a, dev)
    v2, ri2, ro2, ci2 = to_dev(a2, dev)
    bd, bd2 = _t(b, dev), _t(b2, dev)
    lhs, rhs = _t(dense_operand((m, 512), 23, "pm1"), dev), _t(dense_operand((k, 512), 24, "pm1"), dev)
    ops.plan_policy(0)
    want = (ops.spmm(m, k, v, ri, ro, ci, bd), ops.sddmm(m, k, ri, ro, ci, lhs, rhs),
            ops.spmm(m, k, v2, ri2, ro2, ci2, bd2))
    ops.plan_policy(1)
    side = torch.cuda.Stream(dev)
    side.wait_stream(torch.cuda.current_stream(dev))
    with torch.cuda.stream(side):
        ops.spmm(m, k, v, ri, ro, ci, bd)            # plans topology 1; no synchronize before the capture
    torch.cuda.current_stream(dev).wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        got = (ops.spmm(m, k, v, ri, ro, ci, bd), ops.sddmm(m, k, ri, ro, ci, lhs, rhs),
               ops.spmm(m, k, v2, ri2, ro2, ci2, bd2))          # topology 2 is first seen here
    for _ in range(2):
        graph.replay()
    torch.cuda.synchronize()
    for g, w in zip(got, want):
        assert torch.equal(g, w)


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m,n,density,r", [(1024, 1536, 0.1, 5), (700, 300, 0.6, 33), (2048, 2048, 0.05, 1)])
def test_planned_transpose_values_tiled_view_equals_gather(ops, dev, m, n, density, r, dtype):
    """csr_transpose on a cached topology moves the values through the plan's tiled permutation view (large
    inputs) or the plain gather: both equal values[:, perm] exactly, for a last tile that is not full too."""
    a = random_csr(m, n, density, seed=m + n, values="pm1")
    assert a.nnz > 2 * 32768 and a.nnz % 32768 != 0
    ro = torch.from_numpy(a.row_offsets).to(dev)
    ci = torch.from_numpy(a.column_indices).to(dev)
    v = torch.rand((r, a.nnz), device=dev).to(dtype)
    _, _, eperm = oracle.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)[1:]
    ref = v[:, torch.from_numpy(eperm).to(dev).long()]
    try:
        for knob in (2, 1, 0):
            ops.set_variant("permute_tiled", knob)
            vt, rot, cit, perm = ops.csr_transpose_with_perm(m, n, v, ro, ci)
            assert torch.equal(vt, ref), knob
            assert torch.equal(ops.permute_values(v, perm), ref), knob
            assert torch.equal(ops.permute_values(v[0], perm), ref[0]), knob
    finally:
        ops.set_variant("permute_tiled", 0)
