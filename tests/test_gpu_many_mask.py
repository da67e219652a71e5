"""Parity of the ``*_many_mask`` op family (one CSR mask per batch element, h heads per mask)
against the oracle applied mask by mask.  The reference only CALLS these ops
(tests/test_attention_many_masks.py:120-150, tests/transformer/functions.py:20-135); the layout is
the one documented in torch_sputnik_b200/csrc/many_mask.inc."""
import numpy as np
import pytest
import torch

import oracle
from oracle import cpu as ocpu
from oracle.csr import dense_operand, random_csr
from helpers import assert_close_fwd

pytestmark = pytest.mark.gpu


def _masks(b, m, n, densities, seed):
    out = []
    for i in range(b):
        c = random_csr(m, n, densities[i % len(densities)], seed=seed + i, values="pm1", swizzle="heaviest")
        out.append(c)
    return out


def _pack(masks, dev, stacked):
    ri = np.stack([c.row_indices for c in masks]).astype(np.int32)
    ro = np.stack([c.row_offsets for c in masks]).astype(np.int32)
    ci = np.concatenate([c.column_indices for c in masks]).astype(np.int32)
    nnz = torch.tensor([c.column_indices.size for c in masks], dtype=torch.int64)
    if not stacked:
        ri, ro = ri.reshape(-1), ro.reshape(-1)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    return nnz, t(ri), t(ro), t(ci)


def _split(flat, masks, heads):
    """flat ragged values (numpy) -> list over masks of [heads, nnz_i]."""
    out, off = [], 0
    for c in masks:
        z = c.column_indices.size
        out.append(flat[off:off + heads * z].reshape(heads, z))
        off += heads * z
    assert off == flat.size
    return out


@pytest.mark.parametrize("stacked", [True, False])
@pytest.mark.parametrize("b,h,m,k,dens", [
    (3, 2, 96, 64, (0.1, 0.3, 0.05)),       # ragged: every mask another nnz
    (2, 4, 128, 32, (0.2,)),                 # k = 32 rows kernel
    (4, 1, 80, 72, (0.5, 0.02)),             # one head per mask, k not a multiple of 32
])
def test_attention_chain_many_mask(b, h, m, k, dens, stacked):
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    masks = _masks(b, m, m, dens, seed=11)
    nnz, ri, ro, ci = _pack(masks, dev, stacked)
    R = b * h
    q = dense_operand((R, m, k), 1, "pm1")
    kk = dense_operand((R, m, k), 2, "pm1")
    v = dense_operand((R, m, k), 3, "pm1")
    tq, tk, tv = (torch.from_numpy(x).to(dev) for x in (q, kk, v))

    tsp.sddmm_many_mask(b, m, m, nnz, ri, ro, ci, tq, tk)          # (first sight: the prefix table is uploaded and cached)
    before = tsp.launch_count()
    scores = tsp.sddmm_many_mask(b, m, m, nnz, ri, ro, ci, tq, tk)
    assert tsp.launch_count() - before == 1                        # ONE kernel for all b masks x h heads
    uniform = len({c.column_indices.size for c in masks}) == 1
    assert scores.dim() == (2 if uniform else 1)
    before = tsp.launch_count()
    probs = tsp.sparse_softmax_many_mask(b, m, nnz, scores, ri, ro, ci)
    out = tsp.spmm_many_mask(b, m, m, nnz, probs, ri, ro, ci, tv)
    assert tsp.launch_count() - before == 2
    assert out.shape == (R, m, k)
    scaled = tsp.sparse_softmax_many_mask_scaled(b, m, nnz, scores, ri, ro, ci, 0.25)
    torch.testing.assert_close(scaled, tsp.sparse_softmax_many_mask(b, m, nnz, scores * 0.25, ri, ro, ci), rtol=1e-5, atol=1e-7)
    torch.cuda.synchronize()

    s_np = _split(scores.cpu().numpy().reshape(-1), masks, h)
    p_np = _split(probs.cpu().numpy().reshape(-1), masks, h)
    o_np = out.cpu().numpy()
    for i, c in enumerate(masks):
        for j in range(h):
            r = i * h + j
            ref = ocpu.sddmm(m, m, c.row_offsets, c.column_indices, q[r], kk[r])
            sc = ocpu.sddmm_abs(m, m, c.row_offsets, c.column_indices, q[r], kk[r])
            assert_close_fwd(s_np[i][j], ref, sc, what=f"sddmm mask {i} head {j}")
            # softmax / spmm are checked on OUR previous stage's output (stage-wise parity)
            pref = ocpu.csr_softmax(s_np[i][j], c.row_offsets)
            np.testing.assert_allclose(p_np[i][j], pref, rtol=2e-5, atol=1e-7)
            oref = ocpu.spmm(m, m, p_np[i][j], c.row_offsets, c.column_indices, v[r])
            osc = ocpu.spmm_abs(m, m, p_np[i][j], c.row_offsets, c.column_indices, v[r])
            assert_close_fwd(o_np[r], oref, osc, what=f"spmm mask {i} head {j}")


@pytest.mark.parametrize("stacked", [True, False])
def test_csr_transpose_many_mask(stacked):
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    b, h, m, n = 3, 2, 70, 90
    masks = _masks(b, m, n, (0.15, 0.4, 0.01), seed=5)
    nnz, ri, ro, ci = _pack(masks, dev, stacked)
    total = int(nnz.sum())
    vals = dense_operand((h * total,), 9, "pm1")
    vt, rot, cit = tsp.csr_transpose_many_mask(b, m, n, nnz, torch.from_numpy(vals).to(dev), ro, ci)
    before = tsp.launch_count()
    vt2, rot2, cit2 = tsp.csr_transpose_many_mask(b, m, n, nnz, torch.from_numpy(vals).to(dev), ro, ci)
    assert tsp.launch_count() - before == 1                        # every mask sorted once; afterwards one gather launch
    assert torch.equal(vt, vt2) and rot2.data_ptr() == rot.data_ptr() and cit2.data_ptr() == cit.data_ptr()
    torch.cuda.synchronize()
    assert rot.shape == (b, n + 1) and rot.dtype == torch.int32
    v_in = _split(vals, masks, h)
    v_out = _split(vt.cpu().numpy().reshape(-1), masks, h)
    cit_np, off = cit.cpu().numpy(), 0
    for i, c in enumerate(masks):
        z = c.column_indices.size
        for j in range(h):
            rv, rro, rci, _ = ocpu.csr_transpose(m, n, v_in[i][j], c.row_offsets, c.column_indices)
            np.testing.assert_array_equal(v_out[i][j], rv)
            np.testing.assert_array_equal(rot[i].cpu().numpy(), rro)
            np.testing.assert_array_equal(cit_np[off:off + z], rci)
        off += z


def test_many_mask_uniform_matches_shared_mask_ops():
    """b copies of ONE mask must reproduce the shared-mask 3-D ops (same kernels; the replica
    count may move a tile-width choice, so compare to rounding, not bit for bit)."""
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    b, h, m, k = 2, 3, 128, 64
    c = random_csr(m, m, 0.2, seed=3, values="pm1", swizzle="heaviest")
    nnz, ri, ro, ci = _pack([c] * b, dev, True)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)
    ri1, ro1, ci1 = t(c.row_indices), t(c.row_offsets), t(c.column_indices)
    q, kk, v = (t(dense_operand((b * h, m, k), s, "pm1")) for s in (1, 2, 3))
    s_many = tsp.sddmm_many_mask(b, m, m, nnz, ri, ro, ci, q, kk)
    s_one = tsp.sddmm(m, m, ri1, ro1, ci1, q, kk)
    assert s_many.shape == s_one.shape
    torch.testing.assert_close(s_many, s_one, rtol=1e-5, atol=1e-5)
    p_many = tsp.sparse_softmax_many_mask(b, m, nnz, s_many, ri, ro, ci)
    torch.testing.assert_close(p_many, tsp.sparse_softmax(s_many, ri1, ro1, ci1), rtol=1e-6, atol=1e-8)
    o_many = tsp.spmm_many_mask(b, m, m, nnz, p_many, ri, ro, ci, v)
    torch.testing.assert_close(o_many, tsp.spmm(m, m, p_many, ri1, ro1, ci1, v), rtol=1e-5, atol=1e-5)


def test_many_mask_argument_errors():
    import torch_sputnik_b200 as tsp
    dev = torch.device("cuda:0")
    b, h, m, k = 2, 2, 32, 32
    masks = _masks(b, m, m, (0.2, 0.3), seed=1)
    nnz, ri, ro, ci = _pack(masks, dev, True)
    q = torch.zeros(b * h, m, k, device=dev)
    with pytest.raises(RuntimeError):
        tsp.sddmm_many_mask(b, m, m, nnz[:1], ri, ro, ci, q, q)            # one count per mask
    with pytest.raises(RuntimeError):
        tsp.sddmm_many_mask(b, m, m, nnz, ri, ro, ci[:-1], q, q)           # column count mismatch
    with pytest.raises(RuntimeError):
        tsp.sddmm_many_mask(b, m, m, nnz, ri, ro, ci, q[:3], q[:3])        # replicas not b*h
    with pytest.raises(RuntimeError):
        tsp.spmm_many_mask(b, m, m, nnz, torch.zeros(7, device=dev), ri, ro, ci, q)
