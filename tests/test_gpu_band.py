"""GPU: the shared-memory-staged "band" SDDMM / SpMM (kernels/band.cuh; short dense dimension,
attention shapes) against the fp64 oracle -- every storage type, ragged shapes, causal masks
(trimmed block range), empty rows, replicas, unsorted columns (device guard and host verdict)."""
import numpy as np
import pytest
import torch

import oracle
from oracle.csr import dense_operand, from_dense_mask, random_csr

pytestmark = pytest.mark.gpu
DTYPES = [torch.float32, torch.float16, torch.bfloat16]
EPS = {torch.float32: 0.0, torch.float16: 2.0 ** -10, torch.bfloat16: 2.0 ** -7}
TINY = {torch.float32: 1e-30, torch.float16: 2.0 ** -24, torch.bfloat16: 1e-30}


@pytest.fixture(autouse=True)
def _band_forced(ops):
    ops.plan_cache_clear()
    ops.set_variant("spmm", 6)
    ops.set_variant("sddmm", 6)
    yield
    ops.set_variant("spmm", 0)
    ops.set_variant("sddmm", 0)
    ops.plan_policy(1)


def _t(x, dev):
    return torch.from_numpy(np.ascontiguousarray(x)).to(dev)


def _q(x, dtype, dev):
    t = _t(x, dev).to(dtype)
    return t, t.float().cpu().numpy()


def _check(got, ref64, scale, dtype, what):
    got = got.float().cpu().numpy().astype(np.float64)
    bound = EPS[dtype] * np.abs(ref64) + 1e-5 * np.asarray(scale, dtype=np.float64) + TINY[dtype]
    err = np.abs(got - ref64)
    worst = float(np.max(err / bound)) if err.size else 0.0
    assert worst <= 1.0, f"{what}: max err/bound = {worst:.3f} (max abs err {err.max():.3e})"


def _masks():
    rng = np.random.default_rng(0)
    causal = np.tril(rng.random((640, 640)) < 0.3)
    banded = np.zeros((512, 1024), dtype=bool)
    for i in range(512):
        banded[i, max(0, 2 * i - 40):2 * i + 40] = rng.random(min(1024, 2 * i + 40) - max(0, 2 * i - 40)) < 0.6
    holes = rng.random((300, 700)) < 0.15
    holes[5::7] = False                      # empty rows
    holes[:, 256:512] = False                # an empty column block
    return {"causal": causal, "banded": banded, "holes": holes}


# (m, n, d): d is the dense dimension (k of the SDDMM, n of the SpMM)
SHAPES = [(512, 1024, 64, 0.1), (130, 300, 64, 0.3), (256, 512, 32, 0.2), (200, 520, 96, 0.2),
          (192, 384, 128, 0.25), (1024, 2048, 64, 0.02)]


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,n,d,density", SHAPES)
@pytest.mark.parametrize("policy", [1, 0])
def test_sddmm_band(ops, dev, dtype, m, n, d, density, policy):
    if dtype != torch.float32 and d % 64 != 0:
        pytest.skip("16-bit band rows are multiples of 64 elements")
    ops.plan_policy(policy)
    a = random_csr(m, n, density, seed=m + d, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    lhs, lhs32 = _q(dense_operand((3, m, d), 1, "pm1"), dtype, dev)
    rhs, rhs32 = _q(dense_operand((3, n, d), 2, "pm1"), dtype, dev)
    for _ in range(2):          # second call: the plan's verdict is on the host (no guard, no fallback launch)
        out = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
        torch.cuda.synchronize()
    assert out.dtype == dtype and tuple(out.shape) == (3, a.nnz)
    for r in range(3):
        ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs32[r], rhs32[r]).astype(np.float64)
        scale = oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs32[r], rhs32[r])
        _check(out[r], ref, scale, dtype, f"sddmm band {dtype} d={d}")
    if policy == 1:
        before = ops.launch_count()
        ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
        assert ops.launch_count() - before == 1


@pytest.mark.parametrize("dtype", DTYPES)
@pytest.mark.parametrize("m,n,d,density", SHAPES)
@pytest.mark.parametrize("policy", [1, 0])
def test_spmm_band(ops, dev, dtype, m, n, d, density, policy):
    if dtype != torch.float32 and d % 64 != 0:
        pytest.skip("16-bit band rows are multiples of 64 elements")
    ops.plan_policy(policy)
    a = random_csr(m, n, density, seed=n + d, values="pm1", swizzle="heaviest")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    vals, vals32 = _q(dense_operand((2, a.nnz), 3, "pm1"), dtype, dev)
    dense, dense32 = _q(dense_operand((2, n, d), 4, "pm1"), dtype, dev)
    for _ in range(2):
        out = ops.spmm(m, n, vals, ri, ro, ci, dense)
        torch.cuda.synchronize()
    assert out.dtype == dtype and tuple(out.shape) == (2, m, d)
    for r in range(2):
        ref = oracle.spmm(m, n, vals32[r], a.row_offsets, a.column_indices, dense32[r]).astype(np.float64)
        scale = oracle.spmm_abs(m, n, vals32[r], a.row_offsets, a.column_indices, dense32[r])
        _check(out[r], ref, scale, dtype, f"spmm band {dtype} d={d}")
    bias = dense_operand((m,), 5, "pm1")
    got = ops.spmm_bias_relu(m, n, vals[0].contiguous(), ri, ro, ci, _t(bias, dev), dense[0].contiguous())
    ref = oracle.spmm(m, n, vals32[0], a.row_offsets, a.column_indices, dense32[0], bias=bias, relu=True).astype(np.float64)
    scale = oracle.spmm_abs(m, n, vals32[0], a.row_offsets, a.column_indices, dense32[0]) + np.abs(bias)[:, None]
    _check(got, ref, scale, dtype, "spmm band bias relu")
    if policy == 1:
        before = ops.launch_count()
        ops.spmm(m, n, vals, ri, ro, ci, dense)
        assert ops.launch_count() - before == 1


@pytest.mark.parametrize("dtype", [torch.float32, torch.float16])
@pytest.mark.parametrize("name", ["causal", "banded", "holes"])
def test_band_on_structured_masks(ops, dev, dtype, name):
    mask = _masks()[name]
    m, n = mask.shape
    d = 64
    a = from_dense_mask(mask, seed=6, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    lhs, lhs32 = _q(dense_operand((m, d), 7, "pm1"), dtype, dev)
    rhs, rhs32 = _q(dense_operand((n, d), 8, "pm1"), dtype, dev)
    out = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
    ref = oracle.sddmm(m, n, a.row_offsets, a.column_indices, lhs32, rhs32).astype(np.float64)
    _check(out, ref, oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs32, rhs32), dtype, f"sddmm band {name}")
    vals, vals32 = _q(a.values, dtype, dev)
    got = ops.spmm(m, n, vals, ri, ro, ci, rhs)
    ref = oracle.spmm(m, n, vals32, a.row_offsets, a.column_indices, rhs32).astype(np.float64)
    _check(got, ref, oracle.spmm_abs(m, n, vals32, a.row_offsets, a.column_indices, rhs32), dtype, f"spmm band {name}")
    empty = np.flatnonzero(np.diff(a.row_offsets) == 0)
    if empty.size:
        assert float(got[empty].abs().max()) == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_band_declines_unsorted_columns(ops, dev, dtype):
    m, n, d = 256, 512, 64
    a = random_csr(m, n, 0.2, seed=9, values="pm1")
    cols, vals = a.column_indices.copy(), a.values.copy()
    for r in range(m):
        s, e = a.row_offsets[r], a.row_offsets[r + 1]
        cols[s:e] = cols[s:e][::-1]
        vals[s:e] = vals[s:e][::-1]
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(cols, dev)
    lhs, lhs32 = _q(dense_operand((m, d), 10, "pm1"), dtype, dev)
    rhs, rhs32 = _q(dense_operand((n, d), 11, "pm1"), dtype, dev)
    v, v32 = _q(vals, dtype, dev)
    sref = oracle.sddmm(m, n, a.row_offsets, cols, lhs32, rhs32).astype(np.float64)
    sscale = oracle.sddmm_abs(m, n, a.row_offsets, cols, lhs32, rhs32)
    pref = oracle.spmm(m, n, v32, a.row_offsets, cols, rhs32).astype(np.float64)
    pscale = oracle.spmm_abs(m, n, v32, a.row_offsets, cols, rhs32)
    for policy in (0, 1, 1):     # ephemeral (device guard + fallback), planned first call, planned with the verdict
        ops.plan_policy(policy)
        _check(ops.sddmm(m, n, ri, ro, ci, lhs, rhs), sref, sscale, dtype, "sddmm, unsorted")
        _check(ops.spmm(m, n, v, ri, ro, ci, rhs), pref, pscale, dtype, "spmm, unsorted")
        torch.cuda.synchronize()


def test_band_spmm_through_a_value_permutation(ops, dev):
    m, n, d = 384, 512, 64
    a = random_csr(m, n, 0.15, seed=12, values="pm1")
    ri, ro, ci = _t(a.row_indices, dev), _t(a.row_offsets, dev), _t(a.column_indices, dev)
    rot, cit, perm, rit = ops.transposed_topology(m, n, ro, ci)
    vals = _t(dense_operand((2, a.nnz), 13, "pm1"), dev)
    gmat = _t(dense_operand((2, m, d), 14, "pm1"), dev)
    got = ops.spmm_perm(n, m, vals, perm, rit, rot, cit, gmat)
    ops.set_variant("spmm", 1)
    base = ops.spmm_perm(n, m, vals, perm, rit, rot, cit, gmat)
    torch.testing.assert_close(got, base, rtol=1e-5, atol=1e-5)
    dmat = torch.zeros(2, m, n, dtype=torch.float64, device=dev)
    rows = torch.repeat_interleave(torch.arange(m, device=dev), (ro[1:] - ro[:-1]).long())
    dmat[:, rows, ci.long()] = vals.double()
    torch.testing.assert_close(got.double(), dmat.transpose(1, 2) @ gmat.double(), rtol=1e-5, atol=1e-5)
