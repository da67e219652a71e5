"""nn.Modules over the sparse Functions -- mirror of modules/sparse_linear.py:69-89 and
modules/sparse_attention.py:38-128 of the reference (same constructor arguments, attribute
names and tensor layouts), minus the CUDA-only construction-time side effects."""
from __future__ import annotations

import math

import numpy as np
import torch
import torch.nn as nn

from .functional import CsrSoftmax, Sddmm, SparseLinearFunction, Spmm, dense_to_sparse


def generate_mask(m, n, device, sparsity=0.9, round_to=4, seed=None):
    """0/1 mask with int(m*n*sparsity) zeros rounded DOWN to a multiple of 4, shuffled
    (modules/sparse_attention.py:25-36); ``seed`` makes it reproducible."""
    total = m * n
    zeros = int(total * sparsity)
    zeros -= zeros % round_to
    mask = np.ones(total, dtype=np.int64)
    mask[:zeros] = 0
    rng = np.random.default_rng(seed) if seed is not None else np.random
    rng.shuffle(mask)
    return torch.from_numpy(mask).reshape(m, n).to(device)


class SparseLinear(nn.Module):
    """y = W_sparse x.  ``forward`` takes x [B, tokens, in] and -- like the reference
    (modules/sparse_linear.py:85-89) -- returns the TRANSPOSED layout [B, out, tokens]."""

    def __init__(self, input_features, output_features):
        super().__init__()
        self.input_features = input_features
        self.output_features = output_features
        self.weight = nn.Parameter(torch.empty(output_features, input_features))
        self.bias = nn.Parameter(torch.empty(output_features))
        bound = 1.0 / math.sqrt(input_features)
        nn.init.uniform_(self.weight, -bound, bound)
        nn.init.uniform_(self.bias, -bound, bound)

    def setup_sparse_tensors(self):
        values, row_indices, row_offsets, column_indices = dense_to_sparse(self.weight)
        self.values = nn.Parameter(values)
        self.row_indices = row_indices
        self.row_offsets = row_offsets
        self.column_indices = column_indices

    def forward(self, x):
        return self.forward_channels_first(x.transpose(1, 2).contiguous())

    def forward_channels_first(self, xt):
        """xt [B, in, tokens] (contiguous) -> [B, out, tokens]: the layout the kernel works in.  Callers that
        chain sparse layers stay in it and skip the ``transpose().contiguous()`` round trip the reference
        makes around every layer (modules/sparse_linear.py:89, modules/sparse_attention.py:108-126)."""
        return SparseLinearFunction.apply(self.output_features, self.input_features, self.values,
                                          self.row_indices, self.row_offsets, self.column_indices, xt)


class SparseAttention(nn.Module):
    """Multi-head attention with one fixed random sparse mask: SDDMM -> scaled CSR softmax ->
    SpMM over batch x heads replicas (modules/sparse_attention.py:57-98)."""

    def __init__(self, num_heads, embedding_size, max_sequence_length=512, sparsity=0.9,
                 device="cuda", seed=None, linear_sparsity=0.9):
        super().__init__()
        assert embedding_size % num_heads == 0, "Model dimension must be divisible by the number of heads."
        self.head_dim = embedding_size // num_heads
        self.num_heads = num_heads
        self.linears = nn.ModuleList(SparseLinear(embedding_size, embedding_size) for _ in range(4))
        self.m = self.n = max_sequence_length
        self.mask2d = generate_mask(self.m, self.n, device, sparsity=sparsity, seed=seed)
        _, self.row_indices, self.row_offsets, self.column_indices = dense_to_sparse(self.mask2d)
        self.linear_sparsity = linear_sparsity

    def prune_and_setup(self, seed=0):
        """Zero ``linear_sparsity`` of each projection weight and build its CSR tensors."""
        g = torch.Generator(device="cpu").manual_seed(seed)
        for lin in self.linears:
            with torch.no_grad():
                keep = torch.rand(lin.weight.shape, generator=g) >= self.linear_sparsity
                lin.weight.mul_(keep.to(lin.weight.device, lin.weight.dtype))
            lin.setup_sparse_tensors()

    def attention(self, query, key, value, mask=None):
        b, h, s, d = query.shape
        q3d, k3d, v3d = (t.reshape(b * h, s, d).contiguous() for t in (query, key, value))
        scores = Sddmm.apply(self.m, self.n, self.row_indices, self.row_offsets,
                             self.column_indices, q3d, k3d)
        weights = CsrSoftmax.apply(scores, self.row_indices, self.row_offsets,
                                   self.column_indices, 1.0 / math.sqrt(self.head_dim))
        return Spmm.apply(self.m, self.n, weights, self.row_indices, self.row_offsets,
                          self.column_indices, v3d)

    def forward(self, query, key, value, mask=None):
        """Same function as the reference's forward (modules/sparse_attention.py:100-128), with the layout
        round trips fused: a projection's output [B, E, S] IS [B, H, D, S], so one transpose of the last two
        dimensions gives the [B*H, S, D] operand of the sparse ops (the reference copies twice: to [B, S, E]
        and again when flattening the heads); the context goes back to [B, E, S] with one transpose and
        feeds the output projection in its own layout; a shared input (self-attention) is transposed once."""
        batch = query.size(0)
        h, d = self.num_heads, self.head_dim
        cf = {}

        def channels_first(x):
            key_ = id(x)
            if key_ not in cf:
                cf[key_] = x.transpose(1, 2).contiguous()            # [B, E, S]
            return cf[key_]
        q, k, v = (lin.forward_channels_first(channels_first(x)).view(batch, h, d, -1).transpose(2, 3)
                   for lin, x in zip(self.linears, (query, key, value)))   # [B, H, S, D] views
        ctx = self.attention(q, k, v, mask).reshape(batch, h, -1, d)        # [B, H, S, D]
        merged_cf = ctx.transpose(2, 3).reshape(batch, h * d, -1)           # [B, E, S], one copy
        return self.linears[-1].forward_channels_first(merged_cf).transpose(1, 2)
