"""GPU: the REFERENCE'S OWN callers of the hot path, byte for byte, against the shim.

`__graft_entry__.build()` stages modules/{spmm,sddmm,sparse_linear,sparse_attention}.py,
tests/transformer/{functions,utils}.py and tests/test_attention_many_masks.py of
mabdullahsoyturk/Torch-Sputnik unchanged into the git-ignored tests/_ref_callers/ (it travels to the
GPU box with the snapshot, like oracle/_ref).  Here they are imported with
``sys.modules['torch_sputnik']`` = the in-tree extension and run forward + backward at the
reference's own test shapes against dense fp64 autograd.  On a CUDA box a missing staged copy is a
FAILURE, not a skip.  Nothing here exercises torch_sputnik_b200/modules (that is test_gpu_modules.py).
"""
import importlib
import math
import os
import sys

import numpy as np
import pytest
import torch

from helpers import to_dev
from oracle.csr import dense_operand, random_csr, row_of

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
STAGED = os.path.join(HERE, "_ref_callers")
NEEDED = ("modules/spmm.py", "modules/sddmm.py", "modules/sparse_linear.py", "modules/sparse_attention.py",
          "transformer/functions.py", "transformer/utils.py", "attention_many_masks.py")


@pytest.fixture(scope="module")
def ref(ops):
    missing = [p for p in NEEDED if not os.path.exists(os.path.join(STAGED, p))]
    assert not missing, (f"staged reference callers missing under {STAGED}: {missing}; run "
                         "`python __graft_entry__.py` where /root/reference is mounted")
    assert sys.modules["torch_sputnik"] is ops      # what the callers' `import torch_sputnik` resolves to
    for p in (os.path.join(STAGED, "transformer"), STAGED):
        if p not in sys.path:
            sys.path.insert(0, p)

    class Ref:
        spmm = importlib.import_module("modules.spmm")
        sddmm = importlib.import_module("modules.sddmm")
        sparse_linear = importlib.import_module("modules.sparse_linear")
        sparse_attention = importlib.import_module("modules.sparse_attention")
        functions = importlib.import_module("functions")      # tests/transformer/functions.py (`from utils import ...`)
        utils = importlib.import_module("utils")
        many_masks = importlib.import_module("attention_many_masks")
    for mod in (Ref.spmm, Ref.sddmm, Ref.sparse_linear, Ref.sparse_attention, Ref.functions, Ref.utils,
                Ref.many_masks):
        assert os.path.realpath(mod.__file__).startswith(os.path.realpath(STAGED)), mod.__file__
    assert Ref.spmm.torch_sputnik is ops
    return Ref


def _rc(a, dev):
    r = torch.from_numpy(row_of(a.row_offsets).astype(np.int64)).to(dev)
    c = torch.from_numpy(a.column_indices.astype(np.int64)).to(dev)
    return r, c


def _dense_from(a, values, dev):
    r, c = _rc(a, dev)
    out = torch.zeros(*values.shape[:-1], a.m, a.k, dtype=values.dtype, device=dev)
    out[..., r, c] = values
    return out


# ------------------------------------------------------------------------------------------------
# modules/spmm.py:8-74  (shapes: tests/test_spmm.py:13 -> 72x64x72 @ 0.9; tests/test_spmm_3d.py:14 -> r = 256)
@pytest.mark.parametrize("replicas", [0, 256])
def test_reference_spmm_function(ref, dev, replicas):
    m, k, n = 72, 64, 72
    a = random_csr(m, k, 0.1, seed=11, values="pm1", round_to=4)
    _, ri, ro, ci = to_dev(a, dev)
    lead = (replicas,) if replicas else ()
    vals = torch.from_numpy(dense_operand(lead + (a.nnz,), 12, "pm1")).to(dev).requires_grad_(True)
    dense = torch.from_numpy(dense_operand(lead + (k, n), 13, "pm1")).to(dev).requires_grad_(True)
    # the reference's own swizzle expression (modules/spmm.py:4-6): lightest row first, int32
    ri_ref = ref.spmm.diffsort(ro)
    assert ri_ref.dtype == torch.int32 and sorted(ri_ref.tolist()) == list(range(m))
    out = ref.spmm.Spmm.apply(m, k, vals, ri_ref, ro, ci, dense)
    g = torch.from_numpy(dense_operand(tuple(out.shape), 14, "pm1")).to(dev)
    out.backward(g)

    v64, d64 = vals.detach().double().requires_grad_(True), dense.detach().double().requires_grad_(True)
    expect = _dense_from(a, v64, dev) @ d64
    expect.backward(g.double())
    torch.testing.assert_close(out.double(), expect, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dense.grad.double(), d64.grad, rtol=1e-5, atol=1e-5)
    # 3-D backward: the reference hands csr_transpose [R, nnz] values (modules/spmm.py:59), which its own
    # binding cannot take (src/transpose_cuda.cu:58); the shim permutes every replica.
    torch.testing.assert_close(vals.grad.double(), v64.grad, rtol=1e-5, atol=1e-5)
    assert ri is not None


# modules/sddmm.py:9-75  (tests/test_sddmm.py / test_sddmm_3d.py shapes)
@pytest.mark.parametrize("replicas", [0, 64])
def test_reference_sddmm_function(ref, dev, replicas):
    m, k, n = 72, 64, 96
    a = random_csr(m, n, 0.1, seed=21, values="pm1", round_to=4)
    _, ri, ro, ci = to_dev(a, dev)
    lead = (replicas,) if replicas else ()
    lhs = torch.from_numpy(dense_operand(lead + (m, k), 22, "pm1")).to(dev).requires_grad_(True)
    rhs = torch.from_numpy(dense_operand(lead + (n, k), 23, "pm1")).to(dev).requires_grad_(True)
    out = ref.sddmm.Sddmm.apply(m, n, ref.sddmm.diffsort(ro), ro, ci, lhs, rhs)
    g = torch.from_numpy(dense_operand(tuple(out.shape), 24, "pm1")).to(dev)
    out.backward(g)
    l64, r64 = lhs.detach().double().requires_grad_(True), rhs.detach().double().requires_grad_(True)
    r, c = _rc(a, dev)
    expect = (l64 @ r64.transpose(-2, -1))[..., r, c]
    expect.backward(g.double())
    torch.testing.assert_close(out.double(), expect, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(lhs.grad.double(), l64.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(rhs.grad.double(), r64.grad, rtol=1e-5, atol=1e-5)


# modules/sparse_linear.py:18-89  (tests/test_linear_3d.py:100-164: batch 3, W 256x128, 72 tokens;
# tests/test_left_spmm.py:101-154: batch 4, 768 -> 3072, 512 tokens)
@pytest.mark.parametrize("batch,tokens,fin,fout", [(3, 72, 128, 256), (4, 512, 768, 3072)])
def test_reference_sparse_linear(ref, dev, batch, tokens, fin, fout):
    torch.manual_seed(fin)
    lin = ref.sparse_linear.SparseLinear(fin, fout).to(dev)
    w = torch.randn(fout, fin, device=dev) / math.sqrt(fin)
    w = w * (torch.rand_like(w) >= 0.9).float()                 # 90 % pruned
    with torch.no_grad():
        lin.weight.copy_(w)
    lin.setup_sparse_tensors()
    assert lin.row_indices.dtype == torch.int32 and lin.values.numel() == int((w != 0).sum())
    x = torch.randn(batch, tokens, fin, device=dev, requires_grad=True)
    y = lin(x)                                                  # [B, out, tokens]: W @ x^T, no bias (sparse_linear.py:89)
    assert tuple(y.shape) == (batch, fout, tokens)
    g = torch.randn_like(y)
    y.backward(g)                                               # backward: int64 diffsort into left_spmm (sparse_linear.py:57-60)

    w64 = w.double().requires_grad_(True)
    x64 = x.detach().double().requires_grad_(True)
    expect = (x64 @ w64.t()).transpose(1, 2)
    expect.backward(g.double())
    torch.testing.assert_close(y.double(), expect, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(x.grad.double(), x64.grad, rtol=1e-4, atol=1e-4)
    rr, cc = torch.nonzero(w, as_tuple=True)
    # values.grad: [B, nnz] from sddmm, summed over the batch by autograd (sparse_linear.py:44)
    torch.testing.assert_close(lin.values.grad.double(), w64.grad[rr, cc], rtol=1e-4, atol=2e-4)
    # a second step reuses the cached plan / transposed topology: same gradients
    lin.values.grad = None
    x.grad = None
    lin(x).backward(g)
    torch.testing.assert_close(lin.values.grad.double(), w64.grad[rr, cc], rtol=1e-4, atol=2e-4)
    torch.testing.assert_close(x.grad.double(), x64.grad, rtol=1e-4, atol=1e-4)


# modules/sparse_attention.py:38-128 at seq 512 (its default), 8 heads x 64
def test_reference_sparse_attention(ref, dev):
    torch.manual_seed(5)
    np.random.seed(5)                                            # generate_mask shuffles with numpy's global RNG
    heads, emb, seq, batch = 8, 512, 512, 2
    attn = ref.sparse_attention.SparseAttention(heads, emb, max_sequence_length=seq).to(dev)
    assert attn.mask2d.dtype == torch.int64                      # the int64 mask goes through dense_to_sparse (:51-52)
    assert attn.row_indices.dtype == torch.int32 and attn.row_offsets.dtype == torch.int32
    ws = []
    for lin in attn.linears:                                      # weights are torch.empty in the reference: fill + prune
        w = torch.randn(emb, emb, device=dev) / math.sqrt(emb)
        w = w * (torch.rand_like(w) >= 0.9).float() * 3.0
        with torch.no_grad():
            lin.weight.copy_(w)
        lin.setup_sparse_tensors()
        ws.append(w)
    q_in, k_in, v_in = (torch.randn(batch, seq, emb, device=dev, requires_grad=True) for _ in range(3))
    out = attn(q_in, k_in, v_in, None)                            # [B, S, E]
    assert tuple(out.shape) == (batch, seq, emb)
    g = torch.randn_like(out)
    out.backward(g)

    # dense fp64 restatement.  The reference calls torch_sputnik.sparse_softmax directly
    # (sparse_attention.py:76): no autograd node, so NO gradient reaches the scores, q or k -- a
    # reference bug; the probabilities are a constant in backward, mirrored by .detach().
    w64 = [w.double().requires_grad_(True) for w in ws]
    x64 = [t.detach().double().requires_grad_(True) for t in (q_in, k_in, v_in)]
    proj = [(x @ w.t()).view(batch, seq, heads, emb // heads).transpose(1, 2) for x, w in zip(x64, w64)]
    scores = proj[0] @ proj[1].transpose(-2, -1) / math.sqrt(emb // heads)
    scores = scores.masked_fill(attn.mask2d == 0, float("-inf"))
    probs = torch.softmax(scores, dim=-1).detach()
    ctx = (probs @ proj[2]).transpose(1, 2).reshape(batch, seq, emb)
    expect = ctx @ w64[3].t()
    expect.backward(g.double())
    torch.testing.assert_close(out.double(), expect, rtol=2e-4, atol=2e-4)
    torch.testing.assert_close(v_in.grad.double(), x64[2].grad, rtol=2e-4, atol=2e-4)
    for idx in (2, 3):
        rr, cc = torch.nonzero(ws[idx], as_tuple=True)
        torch.testing.assert_close(attn.linears[idx].values.grad.double(), w64[idx].grad[rr, cc], rtol=2e-4, atol=5e-4)
    # the reference bug, pinned: nothing flows back through its softmax
    assert q_in.grad is None and k_in.grad is None
    assert attn.linears[0].values.grad is None and attn.linears[1].values.grad is None


# ------------------------------------------------------------------------------------------------
def _dense_masked_attention(q, k, v, masks, heads, scale):
    """q, k, v [b*h, s, d] fp64; masks [b, s, s] (non-zero = keep)."""
    b = masks.shape[0]
    s, d = q.shape[-2:]
    sc = (q.view(b, heads, s, d) @ k.view(b, heads, s, d).transpose(-2, -1)) * scale
    sc = sc.masked_fill(masks[:, None] == 0, float("-inf"))
    return (torch.softmax(sc, dim=-1) @ v.view(b, heads, s, d)).reshape(b * heads, s, d)


# tests/test_attention_many_masks.py:100-150: the script's own helpers and call sequence
# (flat / concatenated index layout, masks of alternating sparsity 0.2 / 0.5 -> ragged nonzero counts)
def test_reference_many_mask_call_sequence(ref, ops, dev):
    mm = ref.many_masks
    np.random.seed(3)
    torch.manual_seed(3)
    b, n, s, hn = 4, 8, 512, 64
    mask = mm.generate_mask_3d(b, s, s, 0)
    query, key, value = (torch.rand((b, n, s, hn)).cuda() for _ in range(3))
    output_shape = (query.size(2), query.size(0), query.size(1) * query.size(3))
    query, key, value = mm.four_d_to_three_d(query), mm.four_d_to_three_d(key), mm.four_d_to_three_d(value)
    values, row_indices, row_offsets, column_indices, nonzeros = mm.dense_to_sparse_3d(mask)
    assert row_offsets.dim() == 1 and row_offsets.numel() == b * (s + 1)        # concatenated layout
    assert len(set(nonzeros.tolist())) > 1                                      # ragged
    scores = ops.sddmm_many_mask(b, s, s, nonzeros, row_indices, row_offsets, column_indices,
                                 query, key) / math.sqrt(hn)
    weights = ops.sparse_softmax_many_mask(b, s, nonzeros, scores, row_indices, row_offsets, column_indices)
    representations = ops.spmm_many_mask(b, s, s, nonzeros, weights, row_indices, row_offsets,
                                         column_indices, value)
    expect = _dense_masked_attention(query.double(), key.double(), value.double(), mask, n, 1.0 / math.sqrt(hn))
    torch.testing.assert_close(representations.double(), expect, rtol=1e-4, atol=1e-5)
    final = torch.permute(representations, (1, 0, 2)).reshape(*output_shape)
    assert tuple(final.shape) == output_shape


# tests/transformer/functions.py:5-177 -- Spmm / CsrSoftmax / Sddmm over the *_many_mask ops, with the
# stacked layout of tests/transformer/utils.py:20-40; forward and the backward passes the file defines
# (its CsrSoftmax.backward is a placeholder that prints, :105-115 -- not exercised).
def test_reference_transformer_functions(ref, dev):
    fn, ut = ref.functions, ref.utils
    np.random.seed(9)
    torch.manual_seed(9)
    b, h, s, d = 3, 4, 256, 64
    masks = torch.stack([ut.generate_mask(s, s, 0, sparsity=0.9) for _ in range(b)])
    _, row_indices, row_offsets, column_indices, nonzeros = ut.dense_to_sparse_3d(masks)
    assert row_offsets.dim() == 2 and tuple(row_offsets.shape) == (b, s + 1)    # stacked layout
    nnz = int(nonzeros[0])
    lhs = torch.randn(b * h, s, d, device=dev, requires_grad=True)
    rhs = torch.randn(b * h, s, d, device=dev, requires_grad=True)
    val = torch.randn(b * h, s, d, device=dev, requires_grad=True)

    scores = fn.Sddmm.apply(b, s, s, nonzeros, row_indices, row_offsets, column_indices, lhs, rhs)
    assert tuple(scores.shape) == (b * h, nnz)
    probs = fn.CsrSoftmax.apply(b, s, nonzeros, scores.detach() / math.sqrt(d), row_indices, row_offsets,
                                column_indices)
    pv = probs.detach().requires_grad_(True)
    out = fn.Spmm.apply(b, s, s, nonzeros, pv, row_indices, row_offsets, column_indices, val)
    g_out = torch.randn_like(out)
    out.backward(g_out)
    g_sc = torch.randn_like(scores)
    scores.backward(g_sc)

    # dense fp64
    idx = [torch.nonzero(masks[i], as_tuple=True) for i in range(b)]
    l64, r64, v64 = (t.detach().double().requires_grad_(True) for t in (lhs, rhs, val))
    full = l64 @ r64.transpose(-2, -1)
    sc64 = torch.stack([full[i][idx[i // h][0], idx[i // h][1]] for i in range(b * h)])
    torch.testing.assert_close(scores.double(), sc64, rtol=1e-5, atol=1e-4)
    sc64.backward(g_sc.double())
    torch.testing.assert_close(lhs.grad.double(), l64.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(rhs.grad.double(), r64.grad, rtol=1e-4, atol=1e-4)

    dense_sc = torch.full((b * h, s, s), float("-inf"), dtype=torch.float64, device=dev)
    for i in range(b * h):
        dense_sc[i][idx[i // h][0], idx[i // h][1]] = scores.detach().double()[i] / math.sqrt(d)
    p64 = torch.softmax(dense_sc, dim=-1)
    got_p = torch.stack([p64[i][idx[i // h][0], idx[i // h][1]] for i in range(b * h)])
    torch.testing.assert_close(probs.double(), got_p, rtol=1e-5, atol=1e-7)

    pv64 = pv.detach().double().requires_grad_(True)
    dense_p = torch.zeros(b * h, s, s, dtype=torch.float64, device=dev)
    rows = torch.cat([torch.full((nnz,), i, device=dev) for i in range(b * h)])
    r_all = torch.cat([idx[i // h][0] for i in range(b * h)])
    c_all = torch.cat([idx[i // h][1] for i in range(b * h)])
    dense_p = dense_p.index_put((rows, r_all, c_all), pv64.reshape(-1))
    expect = dense_p @ v64
    expect.backward(g_out.double())
    torch.testing.assert_close(out.double(), expect, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(val.grad.double(), v64.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(pv.grad.double(), pv64.grad, rtol=1e-4, atol=1e-4)
