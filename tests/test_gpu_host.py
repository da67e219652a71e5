"""The host-buffer front end (torch_sputnik_b200/host.py): operands in pinned host memory, results
read back to the host, three streams.  Results must equal the oracle's and the plain ops'."""
import numpy as np
import pytest
import torch

from oracle import cpu as ocpu
from oracle.csr import dense_operand, random_csr
from helpers import assert_close_fwd

pytestmark = pytest.mark.gpu


def _pin(a):
    return torch.from_numpy(np.ascontiguousarray(a)).pin_memory()


def test_host_pipeline_spmm_sddmm_many_calls_in_flight():
    import torch_sputnik_b200 as tsp
    from torch_sputnik_b200.host import HostPipeline
    dev = torch.device("cuda:0")
    pipe = HostPipeline(dev, max_inflight=4)
    m, k, n = 256, 192, 160
    cases = []
    for i, dens in enumerate((0.05, 0.3, 0.6, 0.1, 0.25, 0.02)):
        a = random_csr(m, k, dens, seed=40 + i, values="pm1", swizzle="heaviest")
        b = dense_operand((k, n), 100 + i, "pm1")
        lhs = dense_operand((m, n), 200 + i, "pm1")
        rhs = dense_operand((k, n), 300 + i, "pm1")
        cases.append(dict(a=a, b=b, lhs=lhs, rhs=rhs,
                          h={x: _pin(getattr(a, x)) for x in ("values", "row_indices", "row_offsets", "column_indices")},
                          hb=_pin(b), hl=_pin(lhs), hr=_pin(rhs),
                          o1=torch.empty(m, n).pin_memory(), o2=torch.empty(a.nnz).pin_memory()))
    pipe.begin()
    for _ in range(2):                          # second pass overwrites the same host outputs
        for c in cases:
            csr = pipe.upload_csr(c["h"]["values"], c["h"]["row_indices"], c["h"]["row_offsets"],
                                  c["h"]["column_indices"])
            pipe.spmm(m, k, csr, c["hb"], c["o1"])
            pipe.sddmm(m, k, csr, c["hl"], c["hr"], c["o2"])
    pipe.synchronize()
    assert pipe.h2d_bytes > 0 and pipe.d2h_bytes == 2 * sum(c["o1"].numel() * 4 + c["o2"].numel() * 4 for c in cases)
    for c in cases:
        a = c["a"]
        ref = ocpu.spmm(m, k, a.values, a.row_offsets, a.column_indices, c["b"])
        sc = ocpu.spmm_abs(m, k, a.values, a.row_offsets, a.column_indices, c["b"])
        assert_close_fwd(c["o1"].numpy(), ref, sc, what="host spmm")
        ref = ocpu.sddmm(m, k, a.row_offsets, a.column_indices, c["lhs"], c["rhs"])
        sc = ocpu.sddmm_abs(m, k, a.row_offsets, a.column_indices, c["lhs"], c["rhs"])
        assert_close_fwd(c["o2"].numpy(), ref, sc, what="host sddmm")
        # and bit-identical to the device-resident call (same kernels, same launch geometry)
        t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
        d = tsp.spmm(m, k, t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices), t(c["b"]))
        assert torch.equal(d.cpu(), c["o1"])


def test_host_pipeline_softmax_transpose_bias_and_stream_order():
    import torch_sputnik_b200 as tsp  # noqa: F401
    from torch_sputnik_b200.host import HostPipeline
    dev = torch.device("cuda:0")
    pipe = HostPipeline(dev)
    m, n = 100, 130
    a = random_csr(m, n, 0.2, seed=7, values="pm1", swizzle="heaviest")
    h = {x: _pin(getattr(a, x)) for x in ("values", "row_indices", "row_offsets", "column_indices")}
    # host buffers produced by work on the caller's stream: begin() orders the pipeline after it
    side = torch.ones(1 << 22, device=dev).cumsum(0)
    pipe.begin()
    csr = pipe.upload_csr(h["values"], h["row_indices"], h["row_offsets"], h["column_indices"])
    sm = torch.empty(a.nnz).pin_memory()
    pipe.sparse_softmax(csr, sm)
    vt, rot, cit = torch.empty(a.nnz).pin_memory(), torch.empty(n + 1, dtype=torch.int32).pin_memory(), \
        torch.empty(a.nnz, dtype=torch.int32).pin_memory()
    pipe.csr_transpose(m, n, csr, vt, rot, cit)
    bias = dense_operand((m,), 5, "pm1")
    dense = dense_operand((n, 64), 6, "pm1")
    out = torch.empty(m, 64).pin_memory()
    pipe.spmm(m, n, csr, _pin(dense), out, bias=_pin(bias))
    pipe.end()                                   # current stream now waits for the downloads
    torch.cuda.current_stream().synchronize()
    del side
    np.testing.assert_allclose(sm.numpy(), ocpu.csr_softmax(a.values, a.row_offsets), rtol=2e-5, atol=1e-7)
    rv, rro, rci, _ = ocpu.csr_transpose(m, n, a.values, a.row_offsets, a.column_indices)
    np.testing.assert_array_equal(vt.numpy(), rv)
    np.testing.assert_array_equal(rot.numpy(), rro)
    np.testing.assert_array_equal(cit.numpy(), rci)
    ref = ocpu.spmm(m, n, a.values, a.row_offsets, a.column_indices, dense, bias=bias, relu=True)
    sc = ocpu.spmm_abs(m, n, a.values, a.row_offsets, a.column_indices, dense) + np.abs(bias)[:, None]
    assert_close_fwd(out.numpy(), ref, sc, what="host spmm_bias_relu")
