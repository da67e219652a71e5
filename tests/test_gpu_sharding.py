"""GPU: the shards of the paths that shard -- run one after another on ONE device and
reassembled -- equal the unsharded operator.  (tests/test_dist_cpu.py covers the slice arithmetic and
the gloo collectives on CPU; bench.py's `scale_cases` runs the same decomposition across real ranks.)"""
import numpy as np
import pytest
import torch

from helpers import spmm_case, to_dev
from oracle.csr import dense_operand, random_csr
from torch_sputnik_b200 import dist as tdist

pytestmark = pytest.mark.gpu


def _t(x, dev):
    return torch.from_numpy(np.ascontiguousarray(x)).to(dev)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("density,n", [(0.3, 1024), (0.08, 1024), (0.2, 100)])
def test_column_sharded_spmm_reassembles_bit_identically(ops, dev, world, density, n):
    """Every output element is the same fp32 FMA chain over the row's nonzeros in CSR order, whichever
    kernel variant the shard's width selects (gather, staged): bit-identical."""
    m, k = 768, 1024
    a, b = spmm_case(m, k, n, density, seed=world)
    v, ri, ro, ci = to_dev(a, dev)
    bd = _t(b, dev)
    full = ops.spmm(m, k, v, ri, ro, ci, bd)
    parts = []
    for rank in range(world):
        cols = tdist.shard_columns(n, world, rank)
        parts.append(ops.spmm(m, k, v, ri, ro, ci, bd[:, cols].contiguous()))
    assert sum(p.shape[1] for p in parts) == n
    assert torch.equal(torch.cat(parts, dim=1), full)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("density,k", [(0.3, 512), (0.08, 512), (0.1, 64), (0.2, 100)])
def test_row_sharded_sddmm_reassembles(ops, dev, world, density, k):
    """Mask rows cut into nnz-balanced blocks; each shard's output is a contiguous slice of the nnz
    vector.  Bit-identical when the shard takes the same kernel as the full problem; a shard small
    enough to split K over several CTAs sums in a different order -- then within the forward bound."""
    m, n = 1024, 768
    a = random_csr(m, n, density, seed=world + k, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    lhs, rhs = _t(dense_operand((m, k), 1, "pm1"), dev), _t(dense_operand((n, k), 2, "pm1"), dev)
    full = ops.sddmm(m, n, ri, ro, ci, lhs, rhs)
    shards = tdist.shard_rows_by_nnz(ro, world)
    assert shards[0].nnz_begin == 0 and shards[-1].nnz_end == a.nnz
    counts = [s.nnz_end - s.nnz_begin for s in shards]
    assert max(counts) - min(counts) <= 2 * int(np.diff(a.row_offsets).max())  # every cut lands within one row of its target
    parts = []
    for sh in shards:
        m_l, ri_l, ro_l, ci_l = tdist.slice_csr_rows(sh, ro, ci)
        parts.append(ops.sddmm(m_l, n, ri_l, ro_l, ci_l, lhs[sh.row_begin:sh.row_end].contiguous(), rhs))
    got = torch.cat(parts)
    assert got.shape == full.shape
    if not torch.equal(got, full):
        scale = (lhs.abs().double() @ rhs.abs().double().t())[
            torch.repeat_interleave(torch.arange(m, device=dev), (ro[1:] - ro[:-1]).long()), ci.long()]
        assert bool(((got.double() - full.double()).abs() <= 2e-5 * scale + 1e-30).all())


@pytest.mark.parametrize("world", [2, 8])
def test_replica_sharded_attention_chain_reassembles_bit_identically(ops, dev, world):
    """(The full call takes the band SDDMM, a 2-replica shard the rows kernel: same association.)"""
    reps, s, d = 16, 512, 64
    a = random_csr(s, s, 0.1, seed=3, values="pm1")
    _, ri, ro, ci = to_dev(a, dev)
    q, k, v = (_t(dense_operand((reps, s, d), sd, "pm1"), dev) for sd in (4, 5, 6))

    def chain(qq, kk, vv):
        sc = ops.sddmm(s, s, ri, ro, ci, qq, kk)
        return ops.spmm(s, s, ops.sparse_softmax_scaled(sc, ri, ro, ci, 0.125), ri, ro, ci, vv)
    full = chain(q, k, v)
    parts = []
    for rank in range(world):
        sl = tdist.shard_replicas(reps, world, rank)
        parts.append(chain(q[sl].contiguous(), k[sl].contiguous(), v[sl].contiguous()).reshape(-1, s, d))
    assert torch.equal(torch.cat(parts), full)
