"""Host-buffer front end: the torch_sputnik ops called with operands that live in (pinned) HOST
memory, as a user of the reference does when the data comes from a loader
(tests/test_spmm.py:24-30 builds everything on the CPU and ``.cuda()``s it per call).

A call through the plain ops with host data is three serial phases on one stream -- upload,
kernel, download -- and at the benchmark shape (BASELINE.json configs[1]) the PCIe phases are
longer than the kernels.  ``HostPipeline`` runs the three phases on three streams (the copy
engines are independent of the SMs and of each other, PCIe is full duplex), ordered by events
only, so that the upload of call i+1 and the download of call i-1 overlap the kernel of call i.
Nothing here computes: the kernels are the same C-ABI entries, launched on the pipeline's compute
stream through the torch_sputnik module (which launches on torch's current stream).

Memory: device buffers come from torch's caching allocator on the stream that first writes them and
are handed to the consuming stream with ``record_stream`` so that the allocator does not recycle
them early; with a steady workload the allocator reuses the same blocks every step (no cudaMalloc
in the loop).
"""
from __future__ import annotations

import collections

import torch

from . import torch_sputnik as _ops


def bind_to_gpu_numa_node(index: int) -> list[int]:
    """Pin this process to the CPU cores NVML reports as local to GPU `index`, so that the pinned
    host buffers allocated afterwards (first touch) and the copy submission threads live on the
    NUMA node the GPU's PCIe root hangs off.  With one process per GPU on a two-socket box this is
    what keeps eight concurrent uploads from funnelling through one socket.  Best effort: returns
    the cores bound to, or [] when NVML / sched_setaffinity are unavailable."""
    try:
        import os

        import pynvml
        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(h, words)
        cores = [64 * w + b for w, m in enumerate(mask) for b in range(64) if (m >> b) & 1]
        allowed = sorted(set(cores) & os.sched_getaffinity(0))
        if allowed:
            os.sched_setaffinity(0, allowed)
        return allowed
    except Exception:
        return []


class Staged:
    """Device copies of a group of host tensors + the event that marks their arrival."""
    __slots__ = ("tensors", "ready", "nbytes")

    def __init__(self, tensors, ready, nbytes):
        self.tensors, self.ready, self.nbytes = tensors, ready, nbytes

    def __getitem__(self, key):
        return self.tensors[key]


class HostPipeline:
    def __init__(self, device=None, max_inflight: int = 16):
        if not torch.cuda.is_available():
            raise RuntimeError("torch_sputnik_b200.host: no CUDA device (there is no CPU fallback)")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.h2d = torch.cuda.Stream(self.device)
        self.compute = torch.cuda.Stream(self.device)
        self.d2h = torch.cuda.Stream(self.device)
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # device buffers are allocated when a call is ENQUEUED, so an unbounded run-ahead of the
        # host would pile them up; the host waits for the download `max_inflight` calls back
        self.max_inflight = max_inflight
        self._inflight = collections.deque()

    # -- phases ---------------------------------------------------------------------------------
    def begin(self):
        """Order the three streams after the caller's current stream (call before the first upload
        when the host buffers were produced by work on that stream, and for timing)."""
        now = torch.cuda.current_stream(self.device)
        for s in (self.h2d, self.compute, self.d2h):
            s.wait_stream(now)

    def upload(self, **host) -> Staged:
        """Enqueue host->device copies on the upload stream.  Pinned tensors copy asynchronously;
        pageable ones make this call block (cudaMemcpyAsync semantics)."""
        out, nbytes = {}, 0
        with torch.cuda.stream(self.h2d):
            for name, t in host.items():
                d = t.to(self.device, non_blocking=True)
                d.record_stream(self.compute)
                out[name] = d
                nbytes += t.numel() * t.element_size()
            ready = torch.cuda.Event()
            ready.record(self.h2d)
        self.h2d_bytes += nbytes
        return Staged(out, ready, nbytes)

    def call(self, fn, *staged: Staged):
        """Run ``fn()`` (which launches torch_sputnik ops on the current stream) on the compute
        stream once every ``staged`` group has arrived.  Returns (result, done_event)."""
        with torch.cuda.stream(self.compute):
            for s in staged:
                self.compute.wait_event(s.ready)
            result = fn()
            done = torch.cuda.Event()
            done.record(self.compute)
        return result, done

    def download(self, result: torch.Tensor, done, host_out: torch.Tensor):
        """Enqueue the device->host copy of ``result`` into (pinned) ``host_out`` on the download
        stream, after ``done``."""
        result.record_stream(self.d2h)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(done)
            host_out.copy_(result, non_blocking=True)
            landed = torch.cuda.Event()
            landed.record(self.d2h)
        self.d2h_bytes += result.numel() * result.element_size()
        self._inflight.append(landed)
        while len(self._inflight) > self.max_inflight:
            self._inflight.popleft().synchronize()
        return host_out

    def end(self):
        """Make the caller's current stream wait for everything enqueued so far (no host sync)."""
        now = torch.cuda.current_stream(self.device)
        for s in (self.h2d, self.compute, self.d2h):
            now.wait_stream(s)

    def synchronize(self):
        for s in (self.h2d, self.compute, self.d2h):
            s.synchronize()

    # -- the ops, host in / host out --------------------------------------------------------------
    def upload_csr(self, values, row_indices, row_offsets, column_indices) -> Staged:
        kw = dict(row_indices=row_indices, row_offsets=row_offsets, column_indices=column_indices)
        if values is not None:
            kw["values"] = values
        return self.upload(**kw)

    def spmm(self, m, k, csr: Staged, dense: torch.Tensor, out: torch.Tensor, bias=None):
        """out[host] = A @ dense[host] (relu(A @ dense + bias[:, None]) with ``bias``); ``csr`` comes
        from ``upload_csr`` and can be reused by later calls."""
        ops = self.upload(dense=dense, **({"bias": bias} if bias is not None else {}))
        if bias is not None:
            fn = lambda: _ops.spmm_bias_relu(m, k, csr["values"], csr["row_indices"], csr["row_offsets"],
                                             csr["column_indices"], ops["bias"], ops["dense"])
        else:
            fn = lambda: _ops.spmm(m, k, csr["values"], csr["row_indices"], csr["row_offsets"],
                                   csr["column_indices"], ops["dense"])
        res, done = self.call(fn, csr, ops)
        return self.download(res, done, out)

    def sddmm(self, m, n, csr: Staged, lhs: torch.Tensor, rhs: torch.Tensor, out: torch.Tensor):
        ops = self.upload(lhs=lhs, rhs=rhs)
        res, done = self.call(lambda: _ops.sddmm(m, n, csr["row_indices"], csr["row_offsets"],
                                                 csr["column_indices"], ops["lhs"], ops["rhs"]), csr, ops)
        return self.download(res, done, out)

    def sparse_softmax(self, csr: Staged, out: torch.Tensor):
        res, done = self.call(lambda: _ops.sparse_softmax(csr["values"], csr["row_indices"], csr["row_offsets"],
                                                          csr["column_indices"]), csr)
        return self.download(res, done, out)

    def csr_transpose(self, m, n, csr: Staged, values_t, row_offsets_t, column_indices_t):
        res, done = self.call(lambda: _ops.csr_transpose(m, n, csr["values"], csr["row_offsets"],
                                                         csr["column_indices"]), csr)
        for r, h in zip(res, (values_t, row_offsets_t, column_indices_t)):
            self.download(r, done, h)
        return values_t, row_offsets_t, column_indices_t
