"""GPU: the autograd layer (mirror of the reference's modules/) against dense torch autograd."""
import math

import numpy as np
import pytest
import torch

from helpers import to_dev
from oracle.csr import dense_operand, random_csr, row_of

pytestmark = pytest.mark.gpu


def _dense_from(a, values, dev):
    r = torch.from_numpy(row_of(a.row_offsets).astype(np.int64)).to(dev)
    c = torch.from_numpy(a.column_indices.astype(np.int64)).to(dev)
    lead = values.shape[:-1]
    out = torch.zeros(*lead, a.m, a.k, dtype=values.dtype, device=dev)
    out[..., r, c] = values
    return out, r, c


@pytest.mark.parametrize("replicas", [0, 3])
def test_spmm_function_grads(dev, replicas):
    from torch_sputnik_b200.modules import Spmm
    m, k, n = 96, 80, 72
    a = random_csr(m, k, 0.2, seed=1, values="pm1")
    v, ri, ro, ci = to_dev(a, dev)
    lead = (replicas,) if replicas else ()
    vals = (torch.from_numpy(dense_operand(lead + (a.nnz,), 2, "pm1")).to(dev)).requires_grad_(True)
    dense = torch.from_numpy(dense_operand(lead + (k, n), 3, "pm1")).to(dev).requires_grad_(True)
    out = Spmm.apply(m, k, vals, ri, ro, ci, dense)
    g = torch.from_numpy(dense_operand(tuple(out.shape), 4, "pm1")).to(dev)
    out.backward(g)

    vals64 = vals.detach().double().requires_grad_(True)
    dense64 = dense.detach().double().requires_grad_(True)
    A, r, c = _dense_from(a, vals64, dev)
    ref = A @ dense64
    ref.backward(g.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dense.grad.double(), dense64.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(vals.grad.double(), vals64.grad, rtol=1e-5, atol=1e-5)


def test_spmm_function_single_replica_with_2d_values(dev):
    """values [1, nnz] x dense [1, k, n]: the forward returns 2-D (src/spmm_cuda.cu:46) and sddmm returns
    [nnz] (src/sddmm_cuda.cu:43); the backward must still hand autograd a [1, nnz] gradient."""
    from torch_sputnik_b200.modules import Spmm
    m, k, n = 64, 48, 40
    a = random_csr(m, k, 0.25, seed=31, values="pm1")
    v, ri, ro, ci = to_dev(a, dev)
    vals = torch.from_numpy(dense_operand((1, a.nnz), 32, "pm1")).to(dev).requires_grad_(True)
    dense = torch.from_numpy(dense_operand((1, k, n), 33, "pm1")).to(dev).requires_grad_(True)
    out = Spmm.apply(m, k, vals, ri, ro, ci, dense)
    assert tuple(out.shape) == (m, n)
    g = torch.from_numpy(dense_operand((m, n), 34, "pm1")).to(dev)
    out.backward(g)
    assert vals.grad.shape == vals.shape and dense.grad.shape == dense.shape
    vals64 = vals.detach().double().requires_grad_(True)
    dense64 = dense.detach().double().requires_grad_(True)
    A, r, c = _dense_from(a, vals64, dev)
    ref = (A @ dense64)[0]
    ref.backward(g.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(vals.grad.double(), vals64.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(dense.grad.double(), dense64.grad, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("replicas", [0, 3])
def test_sddmm_function_grads(dev, replicas):
    from torch_sputnik_b200.modules import Sddmm
    m, n, k = 96, 112, 64
    a = random_csr(m, n, 0.2, seed=5, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    lead = (replicas,) if replicas else ()
    lhs = torch.from_numpy(dense_operand(lead + (m, k), 6, "pm1")).to(dev).requires_grad_(True)
    rhs = torch.from_numpy(dense_operand(lead + (n, k), 7, "pm1")).to(dev).requires_grad_(True)
    out = Sddmm.apply(m, n, ri, ro, ci, lhs, rhs)
    g = torch.from_numpy(dense_operand(tuple(out.shape), 8, "pm1")).to(dev)
    out.backward(g)
    l64, r64 = lhs.detach().double().requires_grad_(True), rhs.detach().double().requires_grad_(True)
    r = torch.from_numpy(row_of(a.row_offsets).astype(np.int64)).to(dev)
    c = torch.from_numpy(a.column_indices.astype(np.int64)).to(dev)
    ref = (l64 @ r64.transpose(-2, -1))[..., r, c]
    ref.backward(g.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(lhs.grad.double(), l64.grad, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(rhs.grad.double(), r64.grad, rtol=1e-5, atol=1e-5)


def test_sparse_linear_matches_pruned_nn_linear(dev):
    """tests/test_linear_3d.py:100-164: batch 3, W 256x128, 72 tokens; forward and grads within 1e-2
    of a pruned nn.Linear (here 90 % pruned, checked far tighter)."""
    from torch_sputnik_b200.modules import SparseLinear
    torch.manual_seed(0)
    lin = SparseLinear(128, 256).to(dev)
    with torch.no_grad():
        lin.weight.mul_((torch.rand_like(lin.weight) >= 0.9).float())
    lin.setup_sparse_tensors()
    x = torch.randn(3, 72, 128, device=dev, requires_grad=True)
    y = lin(x)                                   # [B, out, tokens] -- transposed, like the reference
    assert tuple(y.shape) == (3, 256, 72)
    g = torch.randn_like(y)
    y.backward(g)
    w64 = lin.weight.detach().double().requires_grad_(True)
    x64 = x.detach().double().requires_grad_(True)
    ref = torch.nn.functional.linear(x64, w64).transpose(1, 2)
    ref.backward(g.double())
    torch.testing.assert_close(y.double(), ref, rtol=1e-5, atol=1e-5)
    torch.testing.assert_close(x.grad.double(), x64.grad, rtol=1e-5, atol=1e-5)
    r = torch.from_numpy(row_of(lin.row_offsets.cpu().numpy()).astype(np.int64)).to(dev)
    c = lin.column_indices.long()
    torch.testing.assert_close(lin.values.grad.double(), w64.grad[r, c], rtol=1e-5, atol=1e-4)


def test_sparse_attention_matches_dense_masked_attention(dev):
    """modules/sparse_attention.py:57-98 pipeline (SDDMM -> scale -> softmax -> SpMM), with a real
    softmax gradient, against dense masked attention in fp64."""
    from torch_sputnik_b200.modules import SparseAttention
    torch.manual_seed(1)
    b, h, s, d = 2, 4, 128, 32
    att = SparseAttention(h, h * d, max_sequence_length=s, sparsity=0.8, device=dev, seed=3).to(dev)
    q, k, v = (torch.randn(b, h, s, d, device=dev, requires_grad=True) for _ in range(3))
    out = att.attention(q, k, v)
    assert tuple(out.shape) == (b * h, s, d)
    g = torch.randn_like(out)
    out.backward(g)
    mask = att.mask2d.bool()
    # no row of the mask may be empty for the dense comparison
    assert bool(mask.any(dim=1).all())
    q64, k64, v64 = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    scores = (q64 @ k64.transpose(-2, -1)) / math.sqrt(d)
    scores = scores.masked_fill(~mask, float("-inf"))
    ref = (torch.softmax(scores, dim=-1) @ v64).reshape(b * h, s, d)
    ref.backward(g.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-5)
    for got, want in ((q.grad, q64.grad), (k.grad, k64.grad), (v.grad, v64.grad)):
        torch.testing.assert_close(got.double(), want, rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("heads", [1, 3])
def test_many_mask_attention_matches_dense_masked_attention(dev, heads):
    """One sparse mask per batch element (the reference's tests/test_attention_many_masks.py
    set-up: alternating sparsities, so every mask has another nnz): forward through the
    ``*_many_mask`` ops, forward + backward through the per-mask autograd Functions, both against
    dense masked attention in fp64."""
    from torch_sputnik_b200.modules import ManyMaskTopology, many_mask_attention
    torch.manual_seed(4)
    b, h, s, d = 3, heads, 96, 64
    masks = torch.stack([torch.rand(s, s, device=dev) >= (0.5 if i % 2 else 0.8) for i in range(b)])
    masks |= torch.eye(s, dtype=torch.bool, device=dev)         # no empty row (dense softmax needs one entry)
    topo = ManyMaskTopology(masks)
    assert len(set(topo.nonzeros.tolist())) > 1
    scale = 1.0 / math.sqrt(d)
    q, k, v = (torch.randn(b, h, s, d, device=dev, requires_grad=True) for _ in range(3))
    out = many_mask_attention(q, k, v, topo, scale)
    g = torch.randn_like(out)
    out.backward(g)
    with torch.no_grad():
        fwd_only = many_mask_attention(q, k, v, topo, scale)     # the *_many_mask op path
    q64, k64, v64 = (t.detach().double().requires_grad_(True) for t in (q, k, v))
    scores = (q64 @ k64.transpose(-2, -1)) * scale
    scores = scores.masked_fill(~masks[:, None], float("-inf"))
    ref = (torch.softmax(scores, dim=-1) @ v64).reshape(b * h, s, d)
    ref.backward(g.double())
    torch.testing.assert_close(out.double(), ref, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(fwd_only.double(), ref, rtol=1e-4, atol=1e-5)
    for got, want in ((q.grad, q64.grad), (k.grad, k64.grad), (v.grad, v64.grad)):
        torch.testing.assert_close(got.double(), want, rtol=1e-4, atol=1e-5)


def test_sparse_attention_module_forward_backward_runs(dev):
    from torch_sputnik_b200.modules import SparseAttention
    torch.manual_seed(2)
    att = SparseAttention(4, 128, max_sequence_length=64, sparsity=0.75, device=dev, seed=4).to(dev)
    att.prune_and_setup(seed=0)
    x = torch.randn(2, 64, 128, device=dev, requires_grad=True)
    y = att(x, x, x)
    assert tuple(y.shape) == (2, 64, 128)
    y.square().mean().backward()
    assert x.grad is not None and torch.isfinite(x.grad).all()
    for lin in att.linears:
        assert lin.values.grad is not None and torch.isfinite(lin.values.grad).all()


def test_sparse_transformer_trains_and_buckets_gradients(dev):
    """configs[4] model at toy size: a few SGD steps through SparseLinear / SparseAttention with the
    gradients accumulated straight into the data-parallel buckets (world size 1 here)."""
    from torch_sputnik_b200 import dist as tdist
    from torch_sputnik_b200.modules.transformer import SparseTransformer
    torch.manual_seed(0)
    model = SparseTransformer(layers=2, hidden=64, heads=2, ffn=128, seq_len=64, vocab=50,
                              linear_sparsity=0.7, attn_sparsity=0.7, device=dev).to(dev)
    model.setup_sparse()
    sparse = tdist.sparse_value_parameters(model)
    assert len(sparse) == 2 * (4 + 2)
    ids = {id(p) for p in sparse}
    dense = [p for p in model.parameters() if p.requires_grad and id(p) not in ids]
    b_sparse, b_dense = tdist.ValueGradBucket(sparse), tdist.ValueGradBucket(dense)
    tokens = torch.randint(0, 50, (2, 64), device=dev)
    target = torch.randint(0, 50, (2, 64), device=dev)
    losses = []
    for _ in range(8):
        b_sparse.zero_()
        b_dense.zero_()
        loss = torch.nn.functional.cross_entropy(model(tokens).reshape(-1, 50), target.reshape(-1))
        loss.backward()
        b_sparse.all_reduce().wait()
        assert sparse[0].grad.data_ptr() == b_sparse.flat.data_ptr()     # still a view of the bucket
        assert torch.isfinite(b_sparse.flat).all() and float(b_sparse.flat.abs().sum()) > 0
        with torch.no_grad():
            for p in sparse + dense:
                p.add_(p.grad, alpha=-0.05)
        losses.append(float(loss))
    assert all(np.isfinite(losses)) and losses[-1] < losses[0]
