"""GPU: the CUDA path against (a) the committed golden vectors from the reference's kernels and
(b) the reference's kernels run live (oracle/_ref/libsputnik_ref.so travels with the snapshot)."""
import os

import numpy as np
import pytest
import torch

import oracle
from oracle import ref_gpu
from oracle.csr import dense_operand, random_csr

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_b200.npz")
needs_golden = pytest.mark.skipif(not os.path.exists(GOLDEN), reason="golden vectors not generated yet")
needs_ref = pytest.mark.skipif(not ref_gpu.available(), reason="oracle/_ref/libsputnik_ref.so not built")


def cases(g, op):
    return sorted({k.split("_")[1] for k in g.files if k.startswith(op + "_")})


@needs_golden
def test_cuda_path_matches_golden(ops, dev):
    g = np.load(GOLDEN)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    for c in cases(g, "spmm"):
        p = f"spmm_{c}_"
        m, k, n = (int(x) for x in g[p + "shape"])
        args = (m, k, t(g[p + "values"]), t(g[p + "row_indices"]), t(g[p + "row_offsets"]), t(g[p + "column_indices"]))
        if p + "bias" in g.files:
            got = ops.spmm_bias_relu(*args, t(g[p + "bias"]), t(g[p + "dense"]))
        else:
            got = ops.spmm(*args, t(g[p + "dense"]))
        scale = oracle.spmm_abs(m, k, g[p + "values"], g[p + "row_offsets"], g[p + "column_indices"], g[p + "dense"])
        if p + "bias" in g.files:
            scale = scale + np.abs(g[p + "bias"])[:, None]
        err = np.abs(got.cpu().numpy().astype(np.float64) - g[p + "out"])
        assert np.all(err <= 2e-5 * scale + 1e-30), (c, err.max())   # both sides carry fp32 rounding
    for c in cases(g, "sddmm"):
        p = f"sddmm_{c}_"
        m, k, n = (int(x) for x in g[p + "shape"])
        got = ops.sddmm(m, n, t(g[p + "row_indices"]), t(g[p + "row_offsets"]), t(g[p + "column_indices"]),
                        t(g[p + "lhs"]), t(g[p + "rhs"]))
        scale = oracle.sddmm_abs(m, n, g[p + "row_offsets"], g[p + "column_indices"], g[p + "lhs"], g[p + "rhs"])
        err = np.abs(got.cpu().numpy().astype(np.float64) - g[p + "out"])
        assert np.all(err <= 2e-5 * scale + 1e-30), (c, err.max())
    for c in cases(g, "softmax"):
        p = f"softmax_{c}_"
        got = ops.sparse_softmax(t(g[p + "values"]), t(g[p + "row_indices"]), t(g[p + "row_offsets"]),
                                 t(g[p + "column_indices"]))
        np.testing.assert_allclose(got.cpu().numpy(), g[p + "out"], rtol=2e-6, atol=1e-37, err_msg=c)
    for c in cases(g, "transpose"):
        p = f"transpose_{c}_"
        m, n = (int(x) for x in g[p + "shape"])
        vt, rot, cit = ops.csr_transpose(m, n, t(g[p + "values"]), t(g[p + "row_offsets"]), t(g[p + "column_indices"]))
        np.testing.assert_array_equal(rot.cpu().numpy(), g[p + "row_offsets_t"], err_msg=c)
        np.testing.assert_array_equal(cit.cpu().numpy(), g[p + "column_indices_t"], err_msg=c)
        np.testing.assert_array_equal(vt.cpu().numpy(), g[p + "values_t"], err_msg=c)


@needs_ref
@pytest.mark.parametrize("m,k,n,density", [(72, 64, 72, 0.1), (1024, 1024, 256, 0.125), (4096, 4096, 512, 0.1),
                                           (4096, 4096, 64, 0.1)])
def test_spmm_vs_live_reference_kernel(ops, dev, m, k, n, density):
    a = random_csr(m, k, density, seed=m + n, values="pm1")
    b = dense_operand((k, n), 1, "pm1")
    t = lambda x: torch.from_numpy(x).to(dev)
    args = (t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(b))
    ref = ref_gpu.spmm(m, k, *args)
    got = ops.spmm(m, k, *args)
    bound = 2e-5 * torch.from_numpy(oracle.spmm_abs(m, k, a.values, a.row_offsets, a.column_indices, b)).to(dev)
    assert bool(((got - ref).abs() <= bound + 1e-30).all())


@needs_ref
@pytest.mark.parametrize("m,k,n,density", [(72, 64, 72, 1.0), (1024, 128, 1024, 0.125), (4096, 64, 4096, 0.1)])
def test_sddmm_vs_live_reference_kernel(ops, dev, m, k, n, density):
    a = random_csr(m, n, density, seed=m + k, values="pm1")
    lhs, rhs = dense_operand((m, k), 2, "pm1"), dense_operand((n, k), 3, "pm1")
    t = lambda x: torch.from_numpy(x).to(dev)
    args = (t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(lhs), t(rhs))
    ref = ref_gpu.sddmm(m, n, *args)
    got = ops.sddmm(m, n, *args)
    bound = 2e-5 * torch.from_numpy(oracle.sddmm_abs(m, n, a.row_offsets, a.column_indices, lhs, rhs)).to(dev)
    assert bool(((got - ref).abs() <= bound + 1e-30).all())


@needs_ref
def test_softmax_and_transpose_vs_live_reference(ops, dev):
    a = random_csr(2048, 4096, 0.1, seed=9, values="pm1")
    t = lambda x: torch.from_numpy(x).to(dev)
    v, ri, ro, ci = t((a.values * 8).astype(np.float32)), t(a.row_indices), t(a.row_offsets), t(a.column_indices)
    torch.testing.assert_close(ops.sparse_softmax(v, ri, ro, ci), ref_gpu.sparse_softmax(v, ri, ro, ci),
                               rtol=2e-6, atol=1e-37)
    vt, rot, cit = ops.csr_transpose(2048, 4096, v, ro, ci)
    rvt, rrot, rcit = ref_gpu.csr_transpose(2048, 4096, v, ro, ci)      # cuSPARSE ALG1
    assert torch.equal(rot, rrot) and torch.equal(cit, rcit) and torch.equal(vt, rvt)
