"""CPU: pin the oracle to the REFERENCE ITSELF -- tests/golden/ref_b200.npz holds outputs of the
reference's own Sputnik kernels (compiled unmodified for sm_100a) and of cuSPARSE csr2csc on a
B200 (tests/golden/make_golden_gpu.py).  The oracle must reproduce them: fp32 within the
forward-error bound, softmax within 1e-5 relative, transpose bit-exactly."""
import os

import numpy as np
import pytest

import oracle

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_b200.npz")
pytestmark = pytest.mark.skipif(not os.path.exists(GOLDEN), reason="golden vectors not generated yet")


@pytest.fixture(scope="module")
def g():
    return np.load(GOLDEN)


def cases(g, op):
    return sorted({k.split("_")[1] for k in g.files if k.startswith(op + "_")})


def test_golden_file_is_from_a_b200(g):
    assert "B200" in str(g["meta"][0])
    assert len(cases(g, "spmm")) >= 8 and len(cases(g, "sddmm")) >= 8
    assert len(cases(g, "softmax")) >= 3 and len(cases(g, "transpose")) >= 6


def test_oracle_spmm_matches_reference_kernels(g):
    for c in cases(g, "spmm"):
        p = f"spmm_{c}_"
        m, k, n = g[p + "shape"]
        bias = g[p + "bias"] if p + "bias" in g.files else None
        ref = oracle.spmm(m, k, g[p + "values"], g[p + "row_offsets"], g[p + "column_indices"], g[p + "dense"],
                          bias=bias, relu=bias is not None)
        scale = oracle.spmm_abs(m, k, g[p + "values"], g[p + "row_offsets"], g[p + "column_indices"], g[p + "dense"])
        if bias is not None:
            scale = scale + np.abs(bias)[:, None]
        err = np.abs(ref.astype(np.float64) - g[p + "out"])
        assert np.all(err <= 1e-5 * scale + 1e-30), (c, err.max())
        assert err.max() < 1e-4      # the reference's own gtest tolerance (spmm_test.cu.cc:151)


def test_oracle_sddmm_matches_reference_kernels(g):
    for c in cases(g, "sddmm"):
        p = f"sddmm_{c}_"
        m, k, n = g[p + "shape"]
        ref = oracle.sddmm(m, n, g[p + "row_offsets"], g[p + "column_indices"], g[p + "lhs"], g[p + "rhs"])
        scale = oracle.sddmm_abs(m, n, g[p + "row_offsets"], g[p + "column_indices"], g[p + "lhs"], g[p + "rhs"])
        err = np.abs(ref.astype(np.float64) - g[p + "out"])
        assert np.all(err <= 1e-5 * scale + 1e-30), (c, err.max())
        assert err.max() < 1e-4      # sddmm_test.cu.cc:141


def test_oracle_softmax_matches_reference_kernel(g):
    for c in cases(g, "softmax"):
        p = f"softmax_{c}_"
        ref = oracle.csr_softmax(g[p + "values"], g[p + "row_offsets"])
        np.testing.assert_allclose(ref, g[p + "out"], rtol=1e-5, atol=1e-37, err_msg=c)
        ref32 = oracle.csr_softmax_f32(g[p + "values"], g[p + "row_offsets"])
        np.testing.assert_allclose(ref32, g[p + "out"], rtol=2e-6, atol=1e-37, err_msg=c)


def test_oracle_transpose_matches_cusparse_bit_exactly(g):
    for c in cases(g, "transpose"):
        p = f"transpose_{c}_"
        m, n = g[p + "shape"]
        vt, rot, cit, _ = oracle.csr_transpose(m, n, g[p + "values"], g[p + "row_offsets"], g[p + "column_indices"])
        np.testing.assert_array_equal(rot, g[p + "row_offsets_t"], err_msg=c)
        np.testing.assert_array_equal(cit, g[p + "column_indices_t"], err_msg=c)
        np.testing.assert_array_equal(vt, g[p + "values_t"], err_msg=c)


GOLDEN16 = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_b200_f16.npz")


@pytest.mark.skipif(not os.path.exists(GOLDEN16), reason="fp16 golden vectors not generated yet")
def test_oracle_matches_reference_half_spmm():
    """ref_b200_f16.npz: sputnik::CudaSpmm(half2 ...) (cuda_spmm.cu.cc:511-570) on a B200.  The fp64
    oracle on the same fp16 inputs must agree within one rounding of the result to fp16 plus the fp32
    accumulation bound; and within the reference's own 5e-2 (spmm_test.cu.cc:207)."""
    g16 = np.load(GOLDEN16)
    assert "B200" in str(g16["meta"][0])
    names = sorted({k.split("_")[1] for k in g16.files if k.startswith("spmmh_")})
    assert len(names) >= 8
    for c in names:
        p = f"spmmh_{c}_"
        m, k, n = g16[p + "shape"]
        v32, d32 = g16[p + "values"].astype(np.float32), g16[p + "dense"].astype(np.float32)
        bias = g16[p + "bias"] if p + "bias" in g16.files else None
        assert np.all(np.diff(g16[p + "row_offsets"]) % 2 == 0)       # cuda_spmm.h:75-113
        ref = oracle.spmm(m, k, v32, g16[p + "row_offsets"], g16[p + "column_indices"], d32, bias=bias,
                          relu=bias is not None).astype(np.float64)
        scale = oracle.spmm_abs(m, k, v32, g16[p + "row_offsets"], g16[p + "column_indices"], d32)
        if bias is not None:
            scale = scale + np.abs(bias)[:, None]
        err = np.abs(ref - g16[p + "out"].astype(np.float64))
        assert np.all(err <= 2.0 ** -10 * np.abs(ref) + 1e-5 * scale + 1e-30), (c, err.max())
        assert err.max() < 5e-2
