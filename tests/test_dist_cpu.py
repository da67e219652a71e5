"""CPU (gloo, world_size 2): the host-side logic of the sharded paths -- slice arithmetic, the
nnz-balanced row cut, and the one-all-reduce-per-step gradient bucket.  The per-shard arithmetic
is done with the oracle here (tests may use it); the point is that the SHARDING reassembles the
unsharded result exactly and the bucket carries the summed gradient."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle
from oracle.csr import dense_operand, random_csr


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(fn, world, *args):
    port = _free_port()
    ctx = mp.get_context("spawn")
    q = ctx.SimpleQueue()
    procs = [ctx.Process(target=_entry, args=(fn, r, world, port, q) + args) for r in range(world)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
    assert all(p.exitcode == 0 for p in procs), [p.exitcode for p in procs]
    out = {}
    while not q.empty():
        r, v = q.get()
        out[r] = v
    return out


def _entry(fn, rank, world, port, q, *args):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.set_num_threads(1)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        q.put((rank, fn(rank, world, *args)))
    finally:
        dist.destroy_process_group()


# ------------------------------------------------------------------------------ pure arithmetic
@pytest.mark.parametrize("total,world,align", [(128, 8, 1), (4096, 3, 4), (10, 4, 4), (0, 2, 1), (7, 8, 1), (72, 5, 4)])
def test_shard_range_partitions_exactly(total, world, align):
    from torch_sputnik_b200.dist import shard_range
    pos = 0
    for r in range(world):
        b, e = shard_range(total, world, r, align)
        assert b == pos and e >= b
        assert b % align == 0 or b == total
        pos = e
    assert pos == total


def test_rows_by_nnz_is_balanced_and_contiguous():
    from torch_sputnik_b200.dist import shard_rows_by_nnz
    a = random_csr(512, 640, 0.1, seed=3)
    ro = torch.from_numpy(a.row_offsets)
    for world in (1, 2, 4, 8):
        shards = shard_rows_by_nnz(ro, world)
        assert shards[0].row_begin == 0 and shards[-1].row_end == a.m
        assert shards[0].nnz_begin == 0 and shards[-1].nnz_end == a.nnz
        for s, t in zip(shards, shards[1:]):
            assert s.row_end == t.row_begin and s.nnz_end == t.nnz_begin
        sizes = [s.nnz_end - s.nnz_begin for s in shards]
        assert max(sizes) - min(sizes) <= 2 * int(np.diff(a.row_offsets).max())


def test_rows_by_nnz_with_empty_and_skewed_rows():
    from torch_sputnik_b200.dist import shard_rows_by_nnz
    ro = torch.tensor([0, 0, 0, 100, 100, 101, 101, 101], dtype=torch.int32)
    shards = shard_rows_by_nnz(ro, 4)
    assert sum(s.nnz_end - s.nnz_begin for s in shards) == 101
    assert shards[-1].row_end == 7 and all(s.row_end >= s.row_begin for s in shards)


# --------------------------------------------------------------------------------- gloo, world 2
def _sharded_ops(rank, world):
    from torch_sputnik_b200.dist import shard_columns, shard_replicas, shard_rows_by_nnz, slice_csr_rows
    m, k, n = 96, 80, 72
    a = random_csr(m, k, 0.2, seed=1, values="pm1")
    b = dense_operand((k, n), 2, "pm1")
    # SpMM sharded over dense columns: A replicated, rank owns B[:, cols] and C[:, cols].
    cols = shard_columns(n, world, rank)
    part = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, np.ascontiguousarray(b[:, cols]))
    parts = [None] * world
    dist.all_gather_object(parts, (cols.start, cols.stop, part))
    full = np.concatenate([p[2] for p in sorted(parts)], axis=1)
    ref = oracle.spmm(m, k, a.values, a.row_offsets, a.column_indices, b)
    ok_spmm = np.array_equal(full, ref)

    # SDDMM sharded over nnz-balanced row blocks: rhs replicated, rank owns out[nnz slice].
    lhs, rhs = dense_operand((m, 64), 3, "pm1"), dense_operand((k, 64), 4, "pm1")
    shard = shard_rows_by_nnz(torch.from_numpy(a.row_offsets), world)[rank]
    ml, ri, ro, ci = slice_csr_rows(shard, torch.from_numpy(a.row_offsets), torch.from_numpy(a.column_indices))
    assert sorted(ri.tolist()) == list(range(ml))
    piece = oracle.sddmm(ml, k, ro.numpy(), ci.numpy(), np.ascontiguousarray(lhs[shard.row_begin:shard.row_end]), rhs)
    pieces = [None] * world
    dist.all_gather_object(pieces, (shard.nnz_begin, piece))
    out = np.concatenate([p[1] for p in sorted(pieces, key=lambda t: t[0])])
    ok_sddmm = np.array_equal(out, oracle.sddmm(m, k, a.row_offsets, a.column_indices, lhs, rhs))

    # Replicated 3-D op sharded over replicas.
    reps = 5
    sl = shard_replicas(reps, world, rank)
    counts = [None] * world
    dist.all_gather_object(counts, (sl.start, sl.stop))
    ok_rep = sorted(counts)[0][0] == 0 and sorted(counts)[-1][1] == reps and \
        all(x[1] == y[0] for x, y in zip(sorted(counts), sorted(counts)[1:]))
    return ok_spmm, ok_sddmm, ok_rep


def test_sharded_spmm_sddmm_reassemble_world2():
    res = _spawn(_sharded_ops, 2)
    assert res == {0: (True, True, True), 1: (True, True, True)}


def _bucket(rank, world):
    from torch_sputnik_b200.dist import ValueGradBucket, broadcast_parameters
    torch.manual_seed(rank)                      # ranks start different ...
    p1, p2 = torch.nn.Parameter(torch.randn(1000)), torch.nn.Parameter(torch.randn(37))
    broadcast_parameters([p1, p2])               # ... and are made equal
    start = float(p1.detach().sum() + p2.detach().sum())
    bucket = ValueGradBucket([p1, p2], average=True)
    assert p1.grad.data_ptr() == bucket.flat.data_ptr()
    # a "weight gradient" written in place, as the SDDMM kernel does through the grad view
    p1.grad.copy_(torch.full((1000,), float(rank + 1)))
    p2.grad.copy_(torch.arange(37, dtype=torch.float32) * (rank + 1))
    bucket.all_reduce().wait()
    want1 = sum(r + 1 for r in range(world)) / world
    ok = bool(torch.allclose(p1.grad, torch.full((1000,), want1))) and \
        bool(torch.allclose(p2.grad, torch.arange(37, dtype=torch.float32) * want1))
    return ok, round(start, 4), bucket.nbytes


def test_value_grad_bucket_allreduce_world2():
    res = _spawn(_bucket, 2)
    assert res[0][0] and res[1][0]
    assert res[0][1] == res[1][1]                # broadcast made the parameters identical
    assert res[0][2] == 4 * 1037


def _dp_step(rank, world):
    """Data-parallel step of a sparse layer written with dense torch on CPU: per-rank token shards,
    value-gradient = dense weight gradient sampled at the mask, one bucket all-reduce; must equal
    the single-process gradient over all tokens."""
    from torch_sputnik_b200.dist import ValueGradBucket, shard_range
    torch.manual_seed(0)
    out_f, in_f, tokens = 24, 16, 40
    mask = torch.rand(out_f, in_f) < 0.3
    w = torch.randn(out_f, in_f) * mask
    x = torch.randn(tokens, in_f)
    rows, cols = mask.nonzero(as_tuple=True)
    values = torch.nn.Parameter(w[rows, cols].clone())
    bucket = ValueGradBucket([values], average=False)
    b, e = shard_range(tokens, world, rank)
    xl = x[b:e]
    dense_w = torch.zeros(out_f, in_f).index_put((rows, cols), values)
    y = xl @ dense_w.t()
    g = torch.autograd.grad(y.square().sum(), values)[0]
    values.grad.copy_(g)
    bucket.all_reduce().wait()
    # single-process reference over all tokens
    v2 = w[rows, cols].clone().requires_grad_(True)
    y2 = x @ torch.zeros(out_f, in_f).index_put((rows, cols), v2).t()
    ref = torch.autograd.grad(y2.square().sum(), v2)[0]
    return bool(torch.allclose(values.grad, ref, rtol=1e-5, atol=1e-5))


def test_data_parallel_value_gradient_world2():
    assert _spawn(_dp_step, 2) == {0: True, 1: True}
