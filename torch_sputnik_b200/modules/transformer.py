"""A small sparse Transformer encoder built only from this package's layers -- the model of
BASELINE.json configs[4]: the topology of the reference's toy Transformer
(tests/transformer/driver.py:5-26: 6 layers, hidden 512, 8 heads, ffn 2048, sequence 512) with
every ``nn.Linear`` replaced by a 90 %-sparse ``SparseLinear`` and attention through the shared-mask
sparse ops (``SparseAttention``).  It exists to exercise the hot path end to end (SpMM forward,
SDDMM weight gradients, transposed SpMM data gradients, CSR softmax forward/backward) and the
data-parallel all-reduce of the nnz-long value gradients (``torch_sputnik_b200.dist``)."""
from __future__ import annotations

import torch
import torch.nn as nn

from .layers import SparseAttention, SparseLinear


def _prune_and_setup(lin: SparseLinear, sparsity: float, gen: torch.Generator):
    with torch.no_grad():
        keep = torch.rand(lin.weight.shape, generator=gen) >= sparsity
        lin.weight.mul_(keep.to(lin.weight.device, lin.weight.dtype))
    lin.setup_sparse_tensors()
    # the dense weight is only the initialiser of `values`; like the reference's forward
    # (modules/sparse_linear.py:85-89) the layer never reads it (or the bias) again
    lin.weight.requires_grad_(False)
    lin.bias.requires_grad_(False)


class SparseMLP(nn.Module):
    def __init__(self, hidden, ffn):
        super().__init__()
        self.up = SparseLinear(hidden, ffn)
        self.down = SparseLinear(ffn, hidden)

    def forward(self, x):                                   # x [B, S, hidden]
        h = torch.relu(self.up(x))                          # [B, ffn, S]: stays in the kernels' layout
        return self.down.forward_channels_first(h).transpose(1, 2)


class SparseEncoderLayer(nn.Module):
    def __init__(self, hidden, heads, ffn, seq_len, attn_sparsity, device, seed):
        super().__init__()
        self.norm1 = nn.LayerNorm(hidden)
        self.attn = SparseAttention(heads, hidden, max_sequence_length=seq_len, sparsity=attn_sparsity,
                                    device=device, seed=seed)
        self.norm2 = nn.LayerNorm(hidden)
        self.mlp = SparseMLP(hidden, ffn)

    def forward(self, x):
        y = self.norm1(x)
        x = x + self.attn(y, y, y)
        return x + self.mlp(self.norm2(x))


class SparseTransformer(nn.Module):
    def __init__(self, layers=6, hidden=512, heads=8, ffn=2048, seq_len=512, vocab=1000,
                 linear_sparsity=0.9, attn_sparsity=0.9, device="cuda", seed=0):
        super().__init__()
        self.embed = nn.Embedding(vocab, hidden)
        self.pos = nn.Parameter(torch.zeros(seq_len, hidden))
        self.layers = nn.ModuleList(
            SparseEncoderLayer(hidden, heads, ffn, seq_len, attn_sparsity, device, seed + i) for i in range(layers))
        self.norm = nn.LayerNorm(hidden)
        self.head = nn.Linear(hidden, vocab)
        self.linear_sparsity = linear_sparsity
        self.seed = seed

    def setup_sparse(self):
        """Call after ``.to(device)``: prune every SparseLinear and build its CSR tensors there."""
        gen = torch.Generator(device="cpu").manual_seed(self.seed)
        for m in self.modules():
            if isinstance(m, SparseLinear):
                _prune_and_setup(m, self.linear_sparsity, gen)
        return self

    def forward(self, tokens):                              # tokens [B, S] int64
        x = self.embed(tokens) + self.pos
        for layer in self.layers:
            x = layer(x)
        return self.head(self.norm(x))
