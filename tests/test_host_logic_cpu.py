"""Host-side logic that needs no GPU: the many-mask packing, the host pipeline's refusal to run
without a device, the NUMA helper's best-effort contract."""
import numpy as np
import pytest
import torch

from oracle.csr import from_dense_mask


def test_many_mask_topology_layout_matches_per_mask_csr():
    from torch_sputnik_b200.modules import ManyMaskTopology
    rng = np.random.default_rng(0)
    b, m, n = 3, 17, 29
    masks = rng.random((b, m, n)) < np.array([0.2, 0.5, 0.05])[:, None, None]
    masks[1, 4] = False                                      # an empty row (tests/transformer/driver.py:11-14)
    topo = ManyMaskTopology(torch.from_numpy(masks))
    assert topo.row_indices.shape == (b, m) and topo.row_indices.dtype == torch.int32
    assert topo.row_offsets.shape == (b, m + 1) and topo.row_offsets.dtype == torch.int32
    assert topo.nonzeros.device.type == "cpu" and topo.nonzeros.dtype == torch.int64
    off = 0
    for i in range(b):
        ref = from_dense_mask(masks[i], seed=1)
        nnz = ref.column_indices.size
        assert int(topo.nonzeros[i]) == nnz
        np.testing.assert_array_equal(topo.row_offsets[i].numpy(), ref.row_offsets)          # LOCAL offsets
        np.testing.assert_array_equal(topo.column_indices[off:off + nnz].numpy(), ref.column_indices)
        ri, ro, ci = topo.mask(i)
        assert sorted(ri.tolist()) == list(range(m))                                         # a permutation
        lens = np.diff(ref.row_offsets)
        assert all(lens[ri[j]] >= lens[ri[j + 1]] for j in range(m - 1))                      # heaviest first
        assert torch.equal(ro, topo.row_offsets[i]) and torch.equal(ci, topo.column_indices[off:off + nnz])
        off += nnz
    assert off == topo.column_indices.numel()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device behaviour")
def test_host_pipeline_needs_a_device():
    from torch_sputnik_b200.host import HostPipeline
    with pytest.raises(RuntimeError, match="no CUDA device"):
        HostPipeline()


def test_numa_binding_is_best_effort():
    from torch_sputnik_b200.host import bind_to_gpu_numa_node
    cores = bind_to_gpu_numa_node(10 ** 6)                    # no such GPU: must not raise
    assert cores == [] or all(isinstance(c, int) for c in cores)
