"""Multi-GPU plumbing for the paths that shard -- one process per GPU,
``torch.distributed`` (NCCL on the GPUs, gloo in the CPU tests).

The reference has no distributed code at all.  What shards naturally:

* replicated 3-D ops (attention): replicas (batch x heads) split across ranks -- no collective;
* large 2-D SpMM: dense-operand COLUMNS split across ranks (A replicated) -- no collective;
* 2-D SDDMM: mask ROWS split across ranks in nnz-balanced blocks (rhs replicated) -- no
  collective, every rank owns a contiguous slice of the nnz output;
* data-parallel sparse linear layers: ONE all-reduce of the nnz-long ``values.grad`` per step
  (``ValueGradBucket``) -- never a dense [out, in] gradient.

Nothing here touches a kernel; the helpers only compute slices and own the gradient bucket, so
they are exercised on CPU with gloo (tests/test_dist_cpu.py) and on GPUs by bench.py / tools.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Iterable, List, Sequence, Tuple

import torch


# ----------------------------------------------------------------------------- shard arithmetic
def shard_range(total: int, world: int, rank: int, align: int = 1) -> Tuple[int, int]:
    """[begin, end) of rank's share of ``total`` items, in units of ``align`` items; earlier ranks
    take the remainder.  Concatenating all ranks' ranges gives exactly [0, total)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError(f"bad rank {rank} for world {world}")
    units = (total + align - 1) // align
    base, extra = divmod(units, world)
    b = rank * base + min(rank, extra)
    e = b + base + (1 if rank < extra else 0)
    return min(b * align, total), min(e * align, total)


def shard_replicas(replicas: int, world: int, rank: int) -> slice:
    """Replica (batch x heads) slice of a rank; the CSR topology is replicated."""
    b, e = shard_range(replicas, world, rank)
    return slice(b, e)


def shard_columns(n: int, world: int, rank: int, align: int = 4) -> slice:
    """Dense-operand column slice of a rank for a column-sharded SpMM; multiples of ``align``
    columns keep the 16-byte vector paths of the kernels."""
    b, e = shard_range(n, world, rank, align)
    return slice(b, e)


@dataclass
class RowShard:
    row_begin: int
    row_end: int
    nnz_begin: int
    nnz_end: int


def shard_rows_by_nnz(row_offsets: torch.Tensor, world: int) -> List[RowShard]:
    """Cut the rows of a CSR mask into ``world`` contiguous blocks with (nearly) equal nonzero
    counts -- the SDDMM sharding: rank r computes out[nnz_begin:nnz_end] from lhs[row_begin:row_end]."""
    ro = row_offsets.detach().to("cpu", torch.int64)
    m = ro.numel() - 1
    nnz = int(ro[-1])
    cuts = [0]
    for r in range(1, world):
        target = nnz * r // world
        # first row whose start offset is >= target, never behind the previous cut
        row = int(torch.searchsorted(ro, torch.tensor(target), right=False))
        cuts.append(max(cuts[-1], min(row, m)))
    cuts.append(m)
    return [RowShard(cuts[r], cuts[r + 1], int(ro[cuts[r]]), int(ro[cuts[r + 1]])) for r in range(world)]


def slice_csr_rows(shard: RowShard, row_offsets: torch.Tensor, column_indices: torch.Tensor):
    """(m_local, row_indices, row_offsets, column_indices) of the rows of one shard, re-based to 0,
    with the heaviest-row-first swizzle the kernels are given elsewhere."""
    ro = row_offsets[shard.row_begin:shard.row_end + 1] - shard.nnz_begin
    ci = column_indices[shard.nnz_begin:shard.nnz_end]
    lens = ro[1:] - ro[:-1]
    ri = torch.argsort(lens, descending=True, stable=True).to(torch.int32)
    return shard.row_end - shard.row_begin, ri, ro.to(torch.int32).contiguous(), ci.contiguous()


# ------------------------------------------------------------------- data-parallel value gradients
class ValueGradBucket:
    """One flat fp32 buffer holding ``values.grad`` of every sparse layer, all-reduced once.

    ``params`` are the nnz-long ``values`` parameters (``SparseLinear.values``).  After
    construction each ``p.grad`` is a VIEW into the bucket, so the SDDMM weight-gradient kernels
    accumulate straight into it and no copy-in/copy-out is needed.  ``all_reduce`` launches the
    collective asynchronously (on a side CUDA stream when the bucket lives on a GPU, so it
    overlaps whatever the main stream does next -- typically the dgrad SpMM of the next layer);
    ``wait`` joins it and applies the 1/world average.  Only nnz values travel: 10x fewer bytes
    than a dense data-parallel all-reduce at 90 % sparsity.
    """

    def __init__(self, params: Iterable[torch.nn.Parameter], average: bool = True, group=None):
        self.params: List[torch.nn.Parameter] = [p for p in params]
        if not self.params:
            raise ValueError("ValueGradBucket needs at least one parameter")
        dev, dt = self.params[0].device, self.params[0].dtype
        sizes = [p.numel() for p in self.params]
        self.flat = torch.zeros(sum(sizes), dtype=dt, device=dev)
        off = 0
        for p, n in zip(self.params, sizes):
            if p.device != dev or p.dtype != dt:
                raise ValueError("all bucketed parameters must share device and dtype")
            p.grad = self.flat[off:off + n].view_as(p)
            off += n
        self.average = average
        self.group = group
        self._work = None
        self._stream = torch.cuda.Stream(device=dev) if dev.type == "cuda" else None
        self._event = None

    @property
    def nbytes(self) -> int:
        return self.flat.numel() * self.flat.element_size()

    def zero_(self):
        self.flat.zero_()

    def all_reduce(self):
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(self.group) == 1:
            self._work = None
            return self
        if self._stream is not None:
            self._stream.wait_stream(torch.cuda.current_stream(self.flat.device))
            with torch.cuda.stream(self._stream):
                dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group)
                if self.average:
                    self.flat.mul_(1.0 / dist.get_world_size(self.group))
                self._event = torch.cuda.Event()
                self._event.record(self._stream)
        else:
            self._work = dist.all_reduce(self.flat, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
        return self

    def wait(self):
        import torch.distributed as dist
        if self._event is not None:
            torch.cuda.current_stream(self.flat.device).wait_event(self._event)
            self._event = None
        elif self._work is not None:
            self._work.wait()
            self._work = None
            if self.average:
                self.flat.mul_(1.0 / dist.get_world_size(self.group))
        return self


def sparse_value_parameters(module: torch.nn.Module) -> List[torch.nn.Parameter]:
    """The ``values`` parameters of every SparseLinear-like submodule (anything that went through
    ``setup_sparse_tensors``), in module order."""
    out = []
    for sub in module.modules():
        v = getattr(sub, "values", None)
        if isinstance(v, torch.nn.Parameter) and hasattr(sub, "row_offsets"):
            out.append(v)
    return out


def broadcast_parameters(params: Sequence[torch.Tensor], src: int = 0, group=None):
    """Make every rank start from rank ``src``'s values (weights are replicated in DP)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        for p in params:
            dist.broadcast(p.data if isinstance(p, torch.nn.Parameter) else p, src=src, group=group)
