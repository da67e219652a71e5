"""Fused sparse attention forward (torch_sputnik.sparse_attention) against
the oracle's three-op chain (sddmm -> softmax(scale) -> spmm, modules/sparse_attention.py:66-82)
and against this library's own chain."""
import numpy as np
import pytest
import torch

from oracle import cpu as ocpu
from oracle.csr import dense_operand, random_csr

pytestmark = pytest.mark.gpu


def _oracle_attention(a, q, k, v, scale):
    m, n = q.shape[0], k.shape[0]
    s = ocpu.sddmm(m, n, a.row_offsets, a.column_indices, q, k)
    p = ocpu.csr_softmax(s, a.row_offsets, scale)
    return ocpu.spmm(m, n, p, a.row_offsets, a.column_indices, v), s


def _dev(a, dev):
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    return t(a.row_indices), t(a.row_offsets), t(a.column_indices)


@pytest.mark.parametrize("m,n,d,dens,reps", [
    (128, 128, 64, 0.1, 3),
    (200, 333, 32, 0.3, 2),      # rectangular, many multi-batch rows (> 32 nonzeros)
    (77, 150, 96, 0.05, 1),      # 2-D operands
    (64, 512, 128, 0.5, 2),      # long rows: 8 batches, online rescaling
    (96, 40, 64, 0.9, 2),
])
def test_fused_attention_matches_oracle_chain(m, n, d, dens, reps):
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    a = random_csr(m, n, dens, seed=m + n, values="pm1", swizzle="heaviest")
    ri, ro, ci = _dev(a, dev)
    shape = lambda rows: (reps, rows, d) if reps > 1 else (rows, d)
    q, k, v = dense_operand(shape(m), 1, "pm1"), dense_operand(shape(n), 2, "pm1"), dense_operand(shape(n), 3, "pm1")
    q = q * 3.0                                  # spread the scores so that the running maximum moves
    scale = 1.0 / np.sqrt(d)
    out, lse = tsp.sparse_attention(m, n, ri, ro, ci, *(torch.from_numpy(x).to(dev) for x in (q, k, v)), scale)
    assert out.shape == q.shape and lse.shape == q.shape[:-1]
    out, lse = out.cpu().numpy().reshape(reps, m, d), lse.cpu().numpy().reshape(reps, m)
    for r in range(reps):
        qr, kr, vr = (x.reshape(reps, -1, d)[r] for x in (q, k, v))
        ref, s = _oracle_attention(a, qr, kr, vr, scale)
        # |out| <= max |v| <= 1: a convex combination; tolerance 1e-5 of that scale (north_star)
        np.testing.assert_allclose(out[r], ref, rtol=1e-5, atol=1e-5)
        for i in range(m):
            seg = s[a.row_offsets[i]:a.row_offsets[i + 1]].astype(np.float64) * scale
            want = -np.inf if seg.size == 0 else np.log(np.exp(seg - seg.max()).sum()) + seg.max()
            if seg.size:
                assert abs(lse[r, i] - want) <= 1e-5 * max(1.0, abs(want))
            else:
                assert lse[r, i] == -np.inf


def test_fused_attention_empty_rows_unsorted_columns_and_chain_equivalence():
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    m, n, d, reps = 130, 260, 64, 4
    rng = np.random.default_rng(0)
    mask = rng.random((m, n)) < 0.2
    mask[5] = False
    mask[m - 1] = False
    mask[17] = True                                # one full row
    from oracle.csr import from_dense_mask
    a = from_dense_mask(mask, seed=1)
    # shuffle the columns inside every row: the kernel must not assume ascending order
    ci = a.column_indices.copy()
    for i in range(m):
        seg = ci[a.row_offsets[i]:a.row_offsets[i + 1]]
        rng.shuffle(seg)
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    ri, ro, cid = t(a.row_indices), t(a.row_offsets), t(ci)
    q, k, v = (t(dense_operand((reps, rows, d), s, "pm1")) for rows, s in ((m, 1), (n, 2), (n, 3)))
    out, lse = tsp.sparse_attention(m, n, ri, ro, cid, q, k, v, 0.125)
    scores = tsp.sddmm(m, n, ri, ro, cid, q, k)
    probs = tsp.sparse_softmax_scaled(scores, ri, ro, cid, 0.125)
    chain = tsp.spmm(m, n, probs, ri, ro, cid, v)
    torch.testing.assert_close(out, chain, rtol=1e-5, atol=1e-5)
    assert torch.all(out[:, 5] == 0) and torch.all(out[:, m - 1] == 0)
    assert torch.all(torch.isinf(lse[:, 5]))


def test_fused_attention_unsupported_head_dim_uses_chain():
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    m, n, d = 64, 64, 40
    a = random_csr(m, n, 0.3, seed=2, values="pm1", swizzle="heaviest")
    ri, ro, ci = _dev(a, dev)
    q, k, v = dense_operand((m, d), 1, "pm1"), dense_operand((n, d), 2, "pm1"), dense_operand((n, d), 3, "pm1")
    out, lse = tsp.sparse_attention(m, n, ri, ro, ci, *(torch.from_numpy(x).to(dev) for x in (q, k, v)), 0.2)
    ref, _ = _oracle_attention(a, q, k, v, 0.2)
    np.testing.assert_allclose(out.cpu().numpy(), ref, rtol=1e-5, atol=1e-5)
    assert lse.numel() == 0


@pytest.mark.parametrize("dtype", [torch.float16, torch.bfloat16])
@pytest.mark.parametrize("m,n,d,dens,reps", [(128, 128, 64, 0.1, 3), (200, 333, 128, 0.3, 2), (64, 512, 256, 0.5, 1),
                                             (77, 150, 192, 0.1, 2)])
def test_fused_attention_16bit_storage(dtype, m, n, d, dens, reps):
    """q / k / v / out in fp16 or bf16, scores / probabilities / sums fp32: against the fp64 chain on the
    same (exactly representable) inputs, within one rounding of the result + the fp32 chain's 1e-5."""
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    a = random_csr(m, n, dens, seed=m + d, values="pm1")
    a_rows = np.diff(a.row_offsets)
    ri, ro, ci = _dev(a, dev)
    q, k, v = (torch.from_numpy(dense_operand((reps, rows, d), sd, "pm1")).to(dev).to(dtype)
               for rows, sd in ((m, 1), (n, 2), (n, 3)))
    scale = 1.0 / np.sqrt(d)
    out, lse = tsp.sparse_attention(m, n, ri, ro, ci, q, k, v, scale)
    assert out.dtype == dtype and lse.dtype == torch.float32 and tuple(lse.shape) == (reps, m)
    eps = 2.0 ** -10 if dtype == torch.float16 else 2.0 ** -7
    for r in range(reps):
        ref, _ = _oracle_attention(a, *(x[r].float().cpu().numpy() for x in (q, k, v)), scale)
        got = out[r].float().cpu().numpy()
        assert np.all(np.abs(got - ref) <= eps * np.abs(ref) + 2e-5), float(np.abs(got - ref).max())
    assert bool(torch.isinf(lse[:, torch.from_numpy(a_rows == 0).to(dev)]).all())
