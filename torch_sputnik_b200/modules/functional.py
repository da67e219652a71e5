"""Autograd layer over the ``torch_sputnik`` ops -- the host-side mirror of the reference's
``modules/`` package (same class names, argument order and meaning), written against the
B200 extension.

Differences from the reference, all on purpose:

* ``Spmm`` / ``Sddmm`` backward work for replicated (3-D) operands: the reference hands
  ``[R, nnz]`` values to a ``csr_transpose`` that only understands ``[nnz]``
  (modules/spmm.py:59, src/transpose_cuda.cu:58).
* the transpose of a static topology is computed once per (row_offsets, column_indices) pair
  and cached as a permutation (``TransposedTopology``); later backward passes only gather.
* ``CsrSoftmax`` is a real autograd Function (the reference calls the op directly and gets no
  gradient, modules/sparse_attention.py:76) and folds the ``1/sqrt(d)`` scale.
* ``diffsort`` returns int32 heaviest-row-first; the reference's expression sorts lightest
  first.  Results never depend on the swizzle.
"""
from __future__ import annotations

import torch

from .. import torch_sputnik as ops


def diffsort(offsets: torch.Tensor) -> torch.Tensor:
    """Row swizzle from row offsets: rows ordered by decreasing nonzero count (int32)."""
    lens = offsets[1:] - offsets[:-1]
    return torch.argsort(lens, descending=True, stable=True).to(torch.int32)


def dense_to_sparse(matrix: torch.Tensor):
    """(values, row_indices, row_offsets, column_indices) of a dense 2-D tensor -- the contract
    of modules/sparse_linear.py:9-16 (float32 values, int32 indices)."""
    csr = matrix.detach().to_sparse_csr()
    values = csr.values().clone().to(torch.float32)
    row_offsets = csr.crow_indices().to(torch.int32).contiguous()
    column_indices = csr.col_indices().to(torch.int32).contiguous()
    return values, diffsort(row_offsets), row_offsets, column_indices


class TransposedTopology:
    """CSC view of a static CSR topology: (row_offsets_t, column_indices_t, row_indices_t, perm).

    Comes from the topology's plan inside the extension (``ops.transposed_topology``): the stable
    sort runs once per topology, the plan cache holds the index tensors it is keyed on (so a freed
    address cannot be mistaken for a live topology) and compares their version counters on every
    lookup (so an in-place edit plans again)."""

    __slots__ = ("row_offsets", "column_indices", "perm", "row_indices")

    def __init__(self, m, n, row_offsets, column_indices):
        self.row_offsets, self.column_indices, self.perm, self.row_indices = ops.transposed_topology(
            m, n, row_offsets, column_indices)

    @classmethod
    def get(cls, m, n, row_offsets, column_indices) -> "TransposedTopology":
        return cls(m, n, row_offsets, column_indices)


class Spmm(torch.autograd.Function):
    """out = CSR(values) @ dense.  Arguments as modules/spmm.py:26-36."""

    @staticmethod
    def forward(ctx, m, k, values, row_indices, row_offsets, column_indices, dense):
        ctx.m, ctx.k = m, k
        ctx.topology = (row_indices, row_offsets, column_indices)
        ctx.save_for_backward(values, dense)
        return ops.spmm(m, k, values, row_indices, row_offsets, column_indices, dense)

    @staticmethod
    def backward(ctx, grad_output):
        m, k = ctx.m, ctx.k
        row_indices, row_offsets, column_indices = ctx.topology
        values, dense = ctx.saved_tensors
        grad_output = grad_output.contiguous()
        if grad_output.dim() != dense.dim():
            # one replica: the forward returned [m, n] for a [1, k, n] operand (src/spmm_cuda.cu:46)
            grad_output = grad_output.reshape(dense.shape[:-2] + grad_output.shape[-2:])
        grad_values = grad_dense = None
        if ctx.needs_input_grad[2]:
            # dA = (dC @ B^T) sampled at the pattern; sddmm wants the right operand as [k, n].
            if values.dim() == 1 and grad_output.dim() == 3:
                grad_values = ops.sddmm_reduce(m, k, row_indices, row_offsets, column_indices, grad_output,
                                               dense.contiguous())
            else:
                grad_values = ops.sddmm(m, k, row_indices, row_offsets, column_indices, grad_output,
                                        dense.contiguous())
        if ctx.needs_input_grad[6]:
            t = TransposedTopology.get(m, k, row_offsets, column_indices)
            # A^T @ dC on the cached transposed topology, values read through the permutation
            grad_dense = ops.spmm_perm(k, m, values.contiguous(), t.perm, t.row_indices, t.row_offsets,
                                       t.column_indices, grad_output)
            grad_dense = grad_dense.reshape(dense.shape)
        if grad_values is not None:
            # one replica: sddmm follows the reference and returns [nnz] for a [1, nnz] input (src/sddmm_cuda.cu:43)
            grad_values = grad_values.reshape(values.shape)
        return None, None, grad_values, None, None, None, grad_dense


class Sddmm(torch.autograd.Function):
    """out[j] = <lhs[row(j)], rhs[col(j)]>.  Arguments as modules/sddmm.py:26-37."""

    @staticmethod
    def forward(ctx, m, n, row_indices, row_offsets, column_indices, lhs_matrix, rhs_matrix):
        ctx.m, ctx.n = m, n
        ctx.topology = (row_indices, row_offsets, column_indices)
        ctx.save_for_backward(lhs_matrix, rhs_matrix)
        return ops.sddmm(m, n, row_indices, row_offsets, column_indices, lhs_matrix, rhs_matrix)

    @staticmethod
    def backward(ctx, grad_output):
        m, n = ctx.m, ctx.n
        row_indices, row_offsets, column_indices = ctx.topology
        lhs, rhs = ctx.saved_tensors
        grad_output = grad_output.contiguous()
        grad_lhs = grad_rhs = None
        if ctx.needs_input_grad[5]:
            grad_lhs = ops.spmm(m, n, grad_output, row_indices, row_offsets, column_indices,
                                rhs.contiguous()).reshape(lhs.shape)
        if ctx.needs_input_grad[6]:
            t = TransposedTopology.get(m, n, row_offsets, column_indices)
            grad_rhs = ops.spmm_perm(n, m, grad_output, t.perm, t.row_indices, t.row_offsets, t.column_indices,
                                     lhs.contiguous()).reshape(rhs.shape)
        return None, None, None, None, None, grad_lhs, grad_rhs


class CsrSoftmax(torch.autograd.Function):
    """Row softmax over CSR values with a fused pre-scale: softmax(scale * x)."""

    @staticmethod
    def forward(ctx, values, row_indices, row_offsets, column_indices, scale=1.0):
        y = ops.sparse_softmax_scaled(values.contiguous(), row_indices, row_offsets, column_indices,
                                      float(scale))
        ctx.topology = (row_indices, row_offsets, column_indices)
        ctx.scale = float(scale)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, grad_output):
        (y,) = ctx.saved_tensors
        row_indices, row_offsets, column_indices = ctx.topology
        dx = ops.sparse_softmax_backward(y, grad_output.contiguous(), row_indices, row_offsets,
                                         column_indices, ctx.scale)
        return dx, None, None, None, None


class SparseLinearFunction(torch.autograd.Function):
    """out[b] = W_csr @ dense[b] with one weight for every batch entry (``left_spmm``).
    Arguments as modules/sparse_linear.py:19-30; ``dense`` is [B, in_features, tokens]."""

    @staticmethod
    def forward(ctx, m, k, values, row_indices, row_offsets, column_indices, dense):
        ctx.m, ctx.k = m, k
        ctx.topology = (row_indices, row_offsets, column_indices)
        ctx.save_for_backward(values, dense)
        return ops.left_spmm(m, k, values, row_indices, row_offsets, column_indices, dense)

    @staticmethod
    def backward(ctx, grad_output):
        m, k = ctx.m, ctx.k
        row_indices, row_offsets, column_indices = ctx.topology
        values, dense = ctx.saved_tensors
        grad_output = grad_output.contiguous()
        grad_values = grad_dense = None
        if ctx.needs_input_grad[2]:
            # one [nnz] vector straight from the kernel: the replica (batch) sum is folded into its k loop
            grad_values = ops.sddmm_reduce(m, k, row_indices, row_offsets, column_indices, grad_output,
                                           dense.contiguous())
        if ctx.needs_input_grad[6]:
            t = TransposedTopology.get(m, k, row_offsets, column_indices)
            # W^T @ dY on the plan's transposed topology, the weights read through the permutation
            # (one launch; no permuted copy of the values)
            grad_dense = ops.spmm_perm(k, m, values.contiguous(), t.perm, t.row_indices, t.row_offsets,
                                       t.column_indices, grad_output).reshape(dense.shape)
        return None, None, grad_values, None, None, None, grad_dense


class ManyMaskTopology:
    """``b`` CSR masks of one shape (one per batch element) packed the way the ``*_many_mask`` ops
    take them (csrc/many_mask.inc; tests/transformer/utils.py:20-40 of the reference builds the same
    from a ``[b, m, n]`` mask): stacked ``row_indices [b, m]`` / ``row_offsets [b, m+1]`` with LOCAL
    offsets, concatenated ``column_indices``, ``nonzeros`` on the CPU."""

    def __init__(self, masks: torch.Tensor):
        assert masks.dim() == 3, "masks must be [b, m, n]"
        self.b, self.m, self.n = masks.shape
        parts = [dense_to_sparse(masks[i].to(torch.float32)) for i in range(self.b)]
        self.row_indices = torch.stack([p[1] for p in parts]).contiguous()
        self.row_offsets = torch.stack([p[2] for p in parts]).contiguous()
        self.column_indices = torch.cat([p[3] for p in parts]).contiguous()
        self.nonzeros = torch.tensor([p[3].numel() for p in parts], dtype=torch.int64)
        self._parts = [(p[1], p[2], p[3]) for p in parts]

    def mask(self, i: int):
        """(row_indices, row_offsets, column_indices) of mask ``i`` for the single-mask ops."""
        return self._parts[i]


def many_mask_attention(query, key, value, topo: ManyMaskTopology, scale: float):
    """softmax(scale * Q K^T restricted to mask[batch]) V with one sparse mask per batch element;
    ``query`` / ``key`` / ``value`` are [B, H, S, D] with B == topo.b.  Returns [B * H, S, D].

    Without autograd this is three ``*_many_mask`` calls.  With autograd every mask goes through
    the single-mask Functions above (same kernels, one launch per mask either way), which carry
    the backward passes -- the reference's many-mask softmax gradient is a placeholder
    (tests/transformer/functions.py:105-115)."""
    b, h, s, d = query.shape
    assert b == topo.b and s == topo.m
    q3, k3, v3 = (t.reshape(b * h, -1, d).contiguous() for t in (query, key, value))
    if not (torch.is_grad_enabled() and (q3.requires_grad or k3.requires_grad or v3.requires_grad)):
        scores = ops.sddmm_many_mask(b, topo.m, topo.n, topo.nonzeros, topo.row_indices, topo.row_offsets,
                                     topo.column_indices, q3, k3)
        probs = ops.sparse_softmax_many_mask_scaled(b, topo.m, topo.nonzeros, scores, topo.row_indices,
                                                    topo.row_offsets, topo.column_indices, float(scale))
        return ops.spmm_many_mask(b, topo.m, topo.n, topo.nonzeros, probs, topo.row_indices, topo.row_offsets,
                                  topo.column_indices, v3)
    outs = []
    for i in range(b):
        ri, ro, ci = topo.mask(i)
        sl = slice(i * h, (i + 1) * h)
        scores = Sddmm.apply(topo.m, topo.n, ri, ro, ci, q3[sl], k3[sl])
        probs = CsrSoftmax.apply(scores, ri, ro, ci, scale)
        out = Spmm.apply(topo.m, topo.n, probs, ri, ro, ci, v3[sl])
        outs.append(out.reshape(h, topo.m, d))
    return torch.cat(outs, dim=0)
