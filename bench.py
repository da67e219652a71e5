#!/usr/bin/env python
"""bench.py -- the headline benchmark of BASELINE.json on B200.

Metric: useful GFLOP/s of CSR SpMM + SDDMM at M=N=K=4096, density 0.05/0.1/0.25/0.5, fp32
(useful FLOPs = 2*nnz*N resp. 2*nnz*K, the convention of Sputnik's own benchmark,
third_party/sputnik/sputnik/spmm/spmm_benchmark.cu.cc:29-33), next to algorithmic HBM GB/s and
the fraction of the measured HBM roofline.

A "step" = one pass of the hot path over the workload: for each of the 4 densities one SpMM and
one SDDMM through the `torch_sputnik` operator API (8 operator calls).  At N > 1 GPUs the headline
`value` is WEAK scaling with no collective: every rank runs its own independent 4096^3 workload
(so N = 1 equals the single-GPU bench).  The configs that SHARD are measured next to it, in the
same line, as `scale_cases` -- each with a FIXED global problem split over the N ranks:
  c2_strong          one global C2 problem, SpMM over dense-operand columns, SDDMM over mask rows
                     (+ a one-off check that the shards equal the unsharded result)
  c4_attention_fwd   128 attention replicas split over the ranks (no collective)
  c3_sparse_linear   sparse linear fwd + wgrad + dgrad over sharded tokens, one overlapped NCCL
                     all-reduce of the nnz-long value gradient
  c5_dp_transformer  data-parallel sparse Transformer training step (weak: fixed batch per GPU)

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

`--impl reference` times the reference's own CPU implementation of the path -- the dense torch
masked matmul its tests use as ground truth (tests/test_spmm.py:9-10, tests/test_sddmm.py:8-13)
-- on the host cores, same config/metric/unit.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

DENSITIES = (0.05, 0.1, 0.25, 0.5)
DIM = 4096
METRIC = "SpMM+SDDMM useful GFLOP/s, 4096^3, density 0.05-0.5 (fp32 CSR)"
UNIT = "GFLOP/s"


def ncu_dram_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel dispatched for (op, density),
    read from profiles/r02_c2_dram_traffic.json -- written by tools/ncu_traffic.py from one
    `ncu --set full` capture per dispatched kernel of THIS build (None = not captured)."""
    try:
        with open(os.path.join(ROOT, "profiles", "r02_c2_dram_traffic.json")) as f:
            rows = json.load(f)["kernels"]
        return {(r["op"], float(r["density"])): float(r["dram_bytes"]) for r in rows}
    except Exception:
        return {}

_RESULT_OUT = None


def protect_stdout():
    """Route fd 1 to stderr for the rest of the process and keep the real stdout for the result
    line only: NCCL writes its version banner to fd 1 from C, which would precede the JSON line."""
    global _RESULT_OUT
    if _RESULT_OUT is None:
        sys.stdout.flush()
        _RESULT_OUT = os.fdopen(os.dup(1), "w")
        os.dup2(2, 1)


def emit_line(text):
    out = _RESULT_OUT or sys.stdout
    out.write(text + "\n")
    out.flush()


def spmm_bytes(m, k, n, nnz):   # every operand once
    return 8 * nnz + 4 * (2 * m + 1) + 4 * k * n + 4 * m * n


def sddmm_bytes(m, k, n, nnz):
    return 4 * nnz + 4 * (2 * m + 1) + 4 * m * k + 4 * n * k + 4 * nnz


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            p = json.load(f)
        return float(p["hbm_gbs"]), "measured", float(p.get("sm_max_mhz", 1965.0))
    except Exception:
        return 6650.0, "fallback", 1965.0


def config_dict(world: int):
    """The `config` of the result line -- the SAME dict from both arms (ours / reference)."""
    return {"workload": "C2: SpMM+SDDMM M=N=K=4096 fp32 CSR, density 0.05/0.1/0.25/0.5 "
                        "(BASELINE.json configs[1]); 8 operator calls per step",
            "densities": list(DENSITIES),
            "sharding": "none" if world == 1 else
                        f"one independent 4096^3 problem per rank x{world} (weak; no collective)",
            "l2": "inputs larger than L2: 8 cases with distinct buffers (~1.7 GB) cycled per step"}


def make_workload(rank: int, dim: int, densities):
    """Seeded synthetic inputs (oracle.csr semantics == tests/connectors.py:47-59), host side."""
    from oracle.csr import dense_operand, random_csr
    cases = []
    for i, d in enumerate(densities):
        a = random_csr(dim, dim, d, seed=1000 * rank + i, values="pm1", swizzle="heaviest")
        cases.append(dict(density=d, csr=a,
                          dense=dense_operand((dim, dim), 50 + 1000 * rank + i, "pm1"),
                          lhs=dense_operand((dim, dim), 60 + 1000 * rank + i, "pm1"),
                          rhs=dense_operand((dim, dim), 70 + 1000 * rank + i, "pm1")))
    return cases


# ------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU while the timed region runs."""

    def __init__(self, index: int, period=0.05):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception:
            self.ok = False

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksEventReasonHwPowerBrakeSlowdown", 0x80): "hw_power_brake_slowdown",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if mask & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._halt.wait(self.period)

    def stop(self):
        self._halt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


# ------------------------------------------------------------------------------------------
def cpu_reference_step(cases, torch, sample_rows=None):
    """One step of the reference's CPU path: dense torch masked matmul (fp32, all host threads)."""
    for c in cases:
        a = c["_dense_a"] if sample_rows is None else c["_dense_a"][:sample_rows]
        torch.matmul(a, c["_dense_b"])                                   # tests/test_spmm.py:9-10
        lhs = c["_lhs"] if sample_rows is None else c["_lhs"][:sample_rows]
        mask = c["_mask"] if sample_rows is None else c["_mask"][:sample_rows]
        torch.matmul(lhs, c["_rhs"].transpose(-2, -1)).mul_(mask)        # tests/test_sddmm.py:8-13


def prepare_cpu_reference(cases, torch):
    for c in cases:
        a = c["csr"]
        dense_a = torch.from_numpy(a.to_dense())
        c["_dense_a"], c["_mask"] = dense_a, (dense_a != 0).to(torch.float32)
        c["_dense_b"], c["_lhs"], c["_rhs"] = (torch.from_numpy(c[k]) for k in ("dense", "lhs", "rhs"))


def useful_flops(cases, dim):
    return sum(2.0 * c["csr"].nnz * dim * 2 for c in cases)     # SpMM + SDDMM per density


def run_reference_arm(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    import torch
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    cases = make_workload(0, DIM, DENSITIES)
    prepare_cpu_reference(cases, torch)
    # Bound the sample: time one full step; if it is slow, keep only the first rows of every case.
    t0 = time.perf_counter()
    cpu_reference_step(cases, torch)
    first = time.perf_counter() - t0
    budget = 150.0 / max(1, args.steps + args.warmup)
    sample_rows = None
    if first > budget:
        sample_rows = max(256, int(DIM * budget / first) // 256 * 256)
    frac = 1.0 if sample_rows is None else sample_rows / DIM
    for _ in range(max(0, args.warmup - 1)):
        cpu_reference_step(cases, torch, sample_rows)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_step(cases, torch, sample_rows)
    dt = (time.perf_counter() - t0) / args.steps
    value = useful_flops(cases, DIM) * frac / dt / 1e9
    sample = ("full workload: dense torch.matmul(A_masked, B) and (lhs @ rhs^T) * mask at 4096^3 for "
              "4 densities" if sample_rows is None else
              f"first {sample_rows} of 4096 output rows of every case (same 8 ops per step)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt * 1e3 / frac,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": config_dict(args.gpus),
        "impl_detail": "reference CPU path: dense torch masked matmul (tests/test_spmm.py:9-10, tests/test_sddmm.py:8-13)",
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    emit_line(json.dumps(line))
    return 0


# ------------------------------------------------------------------------------------------
def run_ours(args):
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL's version banner must not land on stdout
        dist.init_process_group("nccl", device_id=dev)

    import torch_sputnik_b200  # noqa: F401  (registers torch_sputnik; raises if not built)
    import torch_sputnik as ops

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    from torch_sputnik_b200.host import HostPipeline, bind_to_gpu_numa_node
    numa_cores = bind_to_gpu_numa_node(local) if world > 1 else []   # before the pinned allocations below

    cases = make_workload(rank, DIM, DENSITIES)
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a))
    for c in cases:
        a = c["csr"]
        c["host"] = {k: t(v).pin_memory() for k, v in dict(
            values=a.values, row_indices=a.row_indices, row_offsets=a.row_offsets,
            column_indices=a.column_indices, dense=c["dense"], lhs=c["lhs"], rhs=c["rhs"]).items()}
        c["devt"] = {k: v.to(dev) for k, v in c["host"].items()}
        c["host_out_spmm"] = torch.empty((DIM, DIM), dtype=torch.float32).pin_memory()
        c["host_out_sddmm"] = torch.empty((a.nnz,), dtype=torch.float32).pin_memory()

    def step(events=None):
        for i, c in enumerate(cases):
            d = c["devt"]
            if events is not None:
                events[2 * i][0].record()
            c["out_spmm"] = ops.spmm(DIM, DIM, d["values"], d["row_indices"], d["row_offsets"],
                                     d["column_indices"], d["dense"])
            if events is not None:
                events[2 * i][1].record()
                events[2 * i + 1][0].record()
            c["out_sddmm"] = ops.sddmm(DIM, DIM, d["row_indices"], d["row_offsets"], d["column_indices"],
                                       d["lhs"], d["rhs"])
            if events is not None:
                events[2 * i + 1][1].record()

    pipe = HostPipeline(dev)

    def e2e_step():
        """The same 8 operator calls with every operand in pinned host memory and every result
        read back to the host, through the package's host-buffer API: uploads, kernels and
        downloads on three streams ordered by events (torch_sputnik_b200/host.py)."""
        h0, d0 = pipe.h2d_bytes, pipe.d2h_bytes
        for c in cases:
            h = c["host"]
            csr = pipe.upload_csr(h["values"], h["row_indices"], h["row_offsets"], h["column_indices"])
            pipe.spmm(DIM, DIM, csr, h["dense"], c["host_out_spmm"])
            pipe.sddmm(DIM, DIM, csr, h["lhs"], h["rhs"], c["host_out_sddmm"])
        return pipe.h2d_bytes - h0, pipe.d2h_bytes - d0

    def e2e_step_serial():
        """One stream, upload -> kernel -> download per call (what `.cuda()` + op + `.cpu()` does)."""
        for c in cases:
            h = c["host"]
            dv = {k: v.to(dev, non_blocking=True) for k, v in h.items()}
            o1 = ops.spmm(DIM, DIM, dv["values"], dv["row_indices"], dv["row_offsets"], dv["column_indices"],
                          dv["dense"])
            o2 = ops.sddmm(DIM, DIM, dv["row_indices"], dv["row_offsets"], dv["column_indices"], dv["lhs"],
                           dv["rhs"])
            c["host_out_spmm"].copy_(o1, non_blocking=True)
            c["host_out_sddmm"].copy_(o2, non_blocking=True)

    # ---- device-resident timing ------------------------------------------------------------
    for _ in range(max(3, args.warmup)):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    ev = [[(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
           for _ in range(2 * len(cases))] for _ in range(args.steps)]
    launches0 = ops.launch_count()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    start.record()
    for s in range(args.steps):
        step(ev[s])
    stop.record()
    barrier()
    launches = ops.launch_count() - launches0
    clocks = sampler.stop()
    ms_total = start.elapsed_time(stop)
    per_kernel_ms = np.array([[e[0].elapsed_time(e[1]) for e in row] for row in ev])  # [steps, 8]
    kernel_ms = per_kernel_ms.mean(axis=0)

    # ---- end-to-end timing (host buffers, copies inside the timed region) -------------------
    e2e_steps = max(1, min(args.steps, 10))
    for _ in range(2):
        e2e_step_serial()
    barrier()
    es, ee = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    es.record()
    for _ in range(e2e_steps):
        e2e_step_serial()
    ee.record()
    barrier()
    e2e_serial_ms = es.elapsed_time(ee) / e2e_steps

    for _ in range(2):
        e2e_step()
    pipe.synchronize()
    barrier()
    es, ee = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    es.record()
    pipe.begin()
    for _ in range(e2e_steps):
        h2d, d2h = e2e_step()
    pipe.end()
    ee.record()
    barrier()
    e2e_ms = es.elapsed_time(ee) / e2e_steps

    cases_f16 = half_storage_cases(torch, ops, dev, cases) if rank == 0 else None
    host_ceiling = host_copy_ceiling(torch, dist, dev, world, barrier)
    scale = scale_cases(args, torch, dist, ops, dev, world, rank, cases, barrier)

    # ---- max over ranks ---------------------------------------------------------------------
    times = torch.tensor([ms_total, e2e_ms, e2e_serial_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(times, op=dist.ReduceOp.MAX)
    ms_total, e2e_ms, e2e_serial_ms = (float(x) for x in times.cpu())
    ms_per_step = ms_total / args.steps

    flops_rank = useful_flops(cases, DIM)
    value = flops_rank * world / (ms_per_step * 1e-3) / 1e9
    e2e_value = flops_rank * world / (e2e_ms * 1e-3) / 1e9

    if rank == 0:
        peak, peak_kind, sm_max = measured_peaks()
        fp32_peak = 148 * 128 * 2 * sm_max * 1e6 / 1e12   # TFLOP/s, non-tensor fp32
        table = []
        for i, c in enumerate(cases):
            nnz = c["csr"].nnz
            for j, (op, nbytes) in enumerate((("spmm", spmm_bytes(DIM, DIM, DIM, nnz)),
                                              ("sddmm", sddmm_bytes(DIM, DIM, DIM, nnz)))):
                ms = float(kernel_ms[2 * i + j])
                gf = 2.0 * nnz * DIM / (ms * 1e-3) / 1e9
                gbs = nbytes / (ms * 1e-3) / 1e9
                table.append({"op": op, "density": c["density"], "nnz": nnz, "ms": round(ms, 4),
                              "useful_gflops": round(gf, 1), "algorithmic_gbs": round(gbs, 1),
                              "hbm_frac": round(gbs / peak, 4),
                              "fp32_frac": round(gf / 1e3 / fp32_peak, 4), "algorithmic_bytes": nbytes})
        top = max(table, key=lambda r: r["ms"])
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(3, args.warmup), "ms_per_step": ms_per_step, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(world),
            "impl_detail": "torch_sputnik.spmm / torch_sputnik.sddmm (in-tree sm_100a extension)",
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "steps": e2e_steps,
                    "api": "torch_sputnik_b200.host.HostPipeline (pinned host operands; upload / kernel / "
                           "download streams overlapped across calls)",
                    "serial_ms_per_step": e2e_serial_ms, "numa_bound_cores": len(numa_cores),
                    "pcie_gbs": round((h2d + d2h) / (e2e_ms * 1e-3) / 1e9, 1),
                    "host_copy_ceiling": host_ceiling},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": f"{top['op']} d={top['density']}",
                         "achieved": top["algorithmic_gbs"], "peak": peak, "peak_kind": peak_kind,
                         "unit": "GB/s", "frac": top["hbm_frac"],
                         "traffic": ncu_dram_traffic().get((top["op"], float(top["density"]))),
                         "algorithmic_bytes": top["algorithmic_bytes"],
                         "fp32_frac": top["fp32_frac"], "avg_launch_ms": top["ms"],
                         "note": "unstructured CSR at these shapes is bound by shared-memory/L1 gather "
                                 "bandwidth and fp32 issue, not HBM (DESIGN.md); fp32_frac is vs "
                                 f"{fp32_peak:.1f} TFLOP/s non-tensor peak"},
            "cases": table,
            "cases_f16": cases_f16,
            "scale_cases": scale,
        }
        if not args.no_cpu_baseline:
            line["cpu_baseline"] = cpu_baseline_leg(cases, torch)
        if not args.no_ref_kernel:
            ref = reference_kernel_leg(cases, dev, torch)
            if ref is not None:
                line["reference_kernel_same_box"] = ref
        emit_line(json.dumps(line))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def half_storage_cases(torch, ops, dev, cases):
    """The same eight C2 operator calls with fp16 STORAGE (values, dense operands and results in
    torch.float16; every product and sum fp32 -- the tsp_*_f16 entry points).  Reported
    beside the fp32 `cases`, outside the headline: ms per call and useful GFLOP/s."""
    out = []
    try:
        for c in cases:
            d = c["devt"]
            nnz = c["csr"].nnz
            v16, dense16 = d["values"].half(), d["dense"].half()
            lhs16, rhs16 = d["lhs"].half(), d["rhs"].half()
            fns = (("spmm", lambda: ops.spmm(DIM, DIM, v16, d["row_indices"], d["row_offsets"], d["column_indices"], dense16)),
                   ("sddmm", lambda: ops.sddmm(DIM, DIM, d["row_indices"], d["row_offsets"], d["column_indices"], lhs16, rhs16)))
            for op, fn in fns:
                for _ in range(2):
                    fn()
                torch.cuda.synchronize()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(5):
                    fn()
                e.record()
                torch.cuda.synchronize()
                ms = s.elapsed_time(e) / 5
                out.append({"op": op, "density": c["density"], "ms": round(ms, 4),
                            "useful_gflops": round(2.0 * nnz * DIM / (ms * 1e-3) / 1e9, 1)})
            del v16, dense16, lhs16, rhs16
        return out
    except Exception as exc:   # measurement aid only
        return {"unavailable": repr(exc)}


def host_copy_ceiling(torch, dist, dev, world, barrier):
    """What the box's host memory system can move with ALL ranks copying at once: pinned 256 MB
    host -> device and device -> host, concurrently on two streams, max time over ranks.  This is the
    ceiling of the `e2e` leg (926 MB up + 329 MB down per step and rank): at N = 8 the eight uploads
    share one socket's memory controllers, which bounds e2e scaling whatever the kernels do."""
    try:
        n = 64 * 1024 * 1024   # fp32 elements = 256 MB
        h_up, h_dn = torch.empty(n, dtype=torch.float32).pin_memory(), torch.empty(n, dtype=torch.float32).pin_memory()
        d_up, d_dn = torch.empty(n, device=dev), torch.empty(n, device=dev)
        s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
        reps = 4

        def run():
            with torch.cuda.stream(s1):
                for _ in range(reps):
                    d_up.copy_(h_up, non_blocking=True)
            with torch.cuda.stream(s2):
                for _ in range(reps):
                    h_dn.copy_(d_dn, non_blocking=True)
        run()
        torch.cuda.synchronize()
        barrier()
        t0 = time.perf_counter()
        run()
        torch.cuda.synchronize()
        dt = torch.tensor([time.perf_counter() - t0], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(dt, op=dist.ReduceOp.MAX)
        per_rank = 2 * reps * n * 4 / float(dt) / 1e9
        return {"gbs_per_rank_both_directions": round(per_rank, 1), "gbs_aggregate": round(per_rank * world, 1),
                "how": f"{world} rank(s) copying 256 MB pinned buffers H2D and D2H concurrently, x{reps}, max over ranks"}
    except Exception as exc:   # measurement aid only
        return {"unavailable": repr(exc)}


def _timed(torch, dist, dev, world, barrier, fn, reps, warm=2):
    """ms per call of fn(): CUDA events on the current stream, barrier on both sides, max over ranks."""
    for _ in range(warm):
        fn()
    barrier()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    barrier()
    ms = torch.tensor([s.elapsed_time(e) / reps], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms)


def scale_cases(args, torch, dist, ops, dev, world, rank, cases, barrier):
    """The configs of BASELINE.json that SHARD, each as ms per step at this N with a
    FIXED global problem (strong scaling) unless noted, so 1 -> N speed-ups follow from the driver's
    N = 1, 2, 4, 8 lines.  Outside the headline's timed region; a failing case reports its error
    instead of taking the headline down."""
    from torch_sputnik_b200 import dist as tdist
    from oracle.csr import random_csr      # input generator only
    out = {}
    reps = max(3, min(args.steps, 10))
    t = lambda a: torch.from_numpy(np.ascontiguousarray(a)).to(dev)

    # -- C2, ONE global 4096^3 problem: SpMM over dense-operand columns, SDDMM over nnz-balanced mask rows
    try:
        glob = cases if rank == 0 else make_workload(0, DIM, DENSITIES)
        work, identical, within = [], True, True
        for c in glob:
            a = c["csr"]
            d = c["devt"] if "devt" in c else {k: t(v) for k, v in dict(
                values=a.values, row_indices=a.row_indices, row_offsets=a.row_offsets,
                column_indices=a.column_indices, dense=c["dense"], lhs=c["lhs"], rhs=c["rhs"]).items()}
            cols = tdist.shard_columns(DIM, world, rank)
            dense_l = d["dense"][:, cols].contiguous()
            sh = tdist.shard_rows_by_nnz(d["row_offsets"], world)[rank]
            m_l, ri_l, ro_l, ci_l = tdist.slice_csr_rows(sh, d["row_offsets"], d["column_indices"])
            lhs_l = d["lhs"][sh.row_begin:sh.row_end].contiguous()
            work.append((d, dense_l, m_l, ri_l, ro_l, ci_l, lhs_l))
            if world > 1:      # one-off: the shards reassemble to the unsharded result
                full = ops.spmm(DIM, DIM, d["values"], d["row_indices"], d["row_offsets"], d["column_indices"], d["dense"])
                mine = ops.spmm(DIM, DIM, d["values"], d["row_indices"], d["row_offsets"], d["column_indices"], dense_l)
                identical &= bool(torch.equal(full[:, cols], mine))
                within &= bool(torch.allclose(full[:, cols], mine, rtol=1e-4, atol=1e-3))
                fulls = ops.sddmm(DIM, DIM, d["row_indices"], d["row_offsets"], d["column_indices"], d["lhs"], d["rhs"])
                mines = ops.sddmm(m_l, DIM, ri_l, ro_l, ci_l, lhs_l, d["rhs"])
                identical &= bool(torch.equal(fulls[sh.nnz_begin:sh.nnz_end], mines))
                within &= bool(torch.allclose(fulls[sh.nnz_begin:sh.nnz_end], mines, rtol=1e-4, atol=1e-3))
                del full, mine, fulls, mines

        def c2_step():
            for d, dense_l, m_l, ri_l, ro_l, ci_l, lhs_l in work:
                ops.spmm(DIM, DIM, d["values"], d["row_indices"], d["row_offsets"], d["column_indices"], dense_l)
                ops.sddmm(m_l, DIM, ri_l, ro_l, ci_l, lhs_l, d["rhs"])
        ms = _timed(torch, dist, dev, world, barrier, c2_step, reps)
        flags = torch.tensor([int(identical), int(within)], device=dev)
        if world > 1:
            dist.all_reduce(flags, op=dist.ReduceOp.MIN)
        out["c2_strong"] = {"ms_per_step": round(ms, 4), "scaling": "strong",
                            "what": "ONE global 4096^3 problem, 4 densities: SpMM sharded over dense-operand columns, "
                                    "SDDMM over nnz-balanced mask rows; no collective",
                            "useful_gflops": round(useful_flops(glob, DIM) / (ms * 1e-3) / 1e9, 1),
                            "shards_bit_identical_to_unsharded": bool(flags[0]) if world > 1 else None,
                            "shards_within_tolerance": bool(flags[1]) if world > 1 else None}
        del work, glob
    except Exception as exc:
        out["c2_strong"] = {"error": repr(exc)}

    # -- C4: attention forward chain, R = 128 replicas of one 4096^2 mask (90 % sparse), head dim 64
    try:
        s_len, hd, replicas = 4096, 64, 128
        rs = tdist.shard_replicas(replicas, world, rank)
        r_l = rs.stop - rs.start
        a = random_csr(s_len, s_len, 0.1, seed=0, values="pm1", round_to=4)
        ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
        g = torch.Generator(device=dev).manual_seed(10 + rank)
        q, k, v = (torch.rand((r_l, s_len, hd), generator=g, device=dev) * 2 - 1 for _ in range(3))

        def c4_step(q=q, k=k, v=v):
            sc = ops.sddmm(s_len, s_len, ri, ro, ci, q, k)
            pr = ops.sparse_softmax_scaled(sc, ri, ro, ci, 1.0 / 8.0)
            return ops.spmm(s_len, s_len, pr, ri, ro, ci, v)
        ms = _timed(torch, dist, dev, world, barrier, c4_step, reps)
        qh, kh, vh = q.half(), k.half(), v.half()
        ms16 = _timed(torch, dist, dev, world, barrier, lambda: c4_step(qh, kh, vh), reps)
        out["c4_attention_fwd"] = {"ms_per_step": round(ms, 4), "ms_per_step_f16": round(ms16, 4), "scaling": "strong",
                                   "what": "sddmm -> csr softmax (scale fused) -> spmm, 128 replicas (batch 8 x 16 heads) "
                                           "x seq 4096 x head dim 64, mask 90 % sparse; replicas sharded, topology replicated; "
                                           "no collective", "replicas_per_gpu": r_l}
        del q, k, v, qh, kh, vh
    except Exception as exc:
        out["c4_attention_fwd"] = {"error": repr(exc)}

    # -- C3: sparse linear 4096 x 4096 at 90 % sparsity, 8192 tokens as [16, 512]: forward, weight gradient
    #    (replica-summed SDDMM into the gradient bucket, all-reduced over ranks on a side stream while the
    #    data gradient runs), data gradient.  Tokens (batch entries) sharded.
    try:
        m = k = 4096
        tokens, batch = 512, 16
        b0, b1 = tdist.shard_range(batch, world, rank)
        b_l = max(1, b1 - b0)
        a = random_csr(m, k, 0.1, seed=0, values="pm1")
        ri, ro, ci = t(a.row_indices), t(a.row_offsets), t(a.column_indices)
        w = torch.nn.Parameter(t(a.values))
        bucket = tdist.ValueGradBucket([w])
        g = torch.Generator(device=dev).manual_seed(20 + rank)
        x = torch.rand((b_l, k, tokens), generator=g, device=dev) * 2 - 1     # x^T: [B, in, S]
        dy = torch.rand((b_l, m, tokens), generator=g, device=dev) * 2 - 1    # [B, out, S]
        rot, cit, perm, rit = ops.transposed_topology(m, k, ro, ci)

        def c3_step():
            ops.left_spmm(m, k, w.detach(), ri, ro, ci, x)                                   # forward
            bucket.flat.copy_(ops.sddmm_reduce(m, k, ri, ro, ci, dy, x))                     # weight gradient
            bucket.all_reduce()                                                              # side stream
            ops.left_spmm(k, m, ops.permute_values(w.detach(), perm), rit, rot, cit, dy)     # data gradient
            bucket.wait()
        ms = _timed(torch, dist, dev, world, barrier, c3_step, reps)
        out["c3_sparse_linear"] = {"ms_per_step": round(ms, 4), "scaling": "strong",
                                   "what": "left_spmm forward + sddmm_reduce weight gradient + transposed left_spmm data "
                                           "gradient, W 4096x4096 at 90 % sparsity, 8192 tokens ([16, 512]) sharded over "
                                           "batch entries; ONE NCCL all-reduce of the nnz-long value gradient per step, "
                                           "overlapped with the data gradient", "batch_per_gpu": b_l,
                                   "allreduce_bytes": int(bucket.nbytes) if world > 1 else 0}
        del x, dy, bucket, w
    except Exception as exc:
        out["c3_sparse_linear"] = {"error": repr(exc)}

    # -- C5: data-parallel training step of the sparse Transformer (weak: 4 sequences of 512 per GPU)
    try:
        from torch_sputnik_b200.modules.transformer import SparseTransformer
        torch.manual_seed(0)
        vocab, bsz, seq = 1000, 4, 512
        model = SparseTransformer(6, 512, 8, 2048, seq, vocab, device=dev).to(dev)
        model.setup_sparse()
        sparse = tdist.sparse_value_parameters(model)
        ids = {id(p) for p in sparse}
        dense = [p for p in model.parameters() if p.requires_grad and id(p) not in ids]
        tdist.broadcast_parameters(sparse + dense)
        b_sparse, b_dense = tdist.ValueGradBucket(sparse), tdist.ValueGradBucket(dense)
        gen = torch.Generator(device="cpu").manual_seed(100 + rank)
        tok = torch.randint(0, vocab, (bsz, seq), generator=gen).to(dev)
        tgt = torch.randint(0, vocab, (bsz, seq), generator=gen).to(dev)
        loss_fn = torch.nn.CrossEntropyLoss()

        def c5_step():
            b_sparse.zero_()
            b_dense.zero_()
            loss = loss_fn(model(tok).reshape(-1, vocab), tgt.reshape(-1))
            loss.backward()
            b_sparse.all_reduce()
            b_dense.all_reduce()
            b_sparse.wait()
            b_dense.wait()
            with torch.no_grad():
                for p in sparse + dense:
                    p.add_(p.grad, alpha=-1e-2)
        ms = _timed(torch, dist, dev, world, barrier, c5_step, reps, warm=3)
        out["c5_dp_transformer"] = {"ms_per_step": round(ms, 4), "scaling": "weak",
                                    "tokens_per_s": round(world * bsz * seq / (ms * 1e-3), 1),
                                    "what": "sparse Transformer (6 layers, hidden 512, 8 heads, ffn 2048, seq 512; 90 %-sparse "
                                            "linear layers + shared-mask sparse attention), 4 sequences per GPU, eager "
                                            "launches; one NCCL all-reduce of the nnz-long value gradients + one of the "
                                            "small dense parameters per step",
                                    "allreduce_bytes": int(b_sparse.nbytes + b_dense.nbytes) if world > 1 else 0}
        del model, b_sparse, b_dense
    except Exception as exc:
        out["c5_dp_transformer"] = {"error": repr(exc)}
    torch.cuda.empty_cache()
    return out


def cpu_baseline_leg(cases, torch):
    """The reference's dense torch CPU path on this box's host cores, one bounded step."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    prepare_cpu_reference(cases, torch)
    sample_rows = 1024
    cpu_reference_step(cases, torch, 256)      # warm-up
    t0 = time.perf_counter()
    cpu_reference_step(cases, torch, sample_rows)
    dt = time.perf_counter() - t0
    reps = 1
    while dt < 5.0 and reps < 8:
        cpu_reference_step(cases, torch, sample_rows)
        reps += 1
        dt = time.perf_counter() - t0
    dt /= reps
    value = useful_flops(cases, DIM) * (sample_rows / DIM) / dt / 1e9
    return {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"first {sample_rows} of 4096 output rows of all 8 ops, {reps} repetition(s); dense torch "
                      "masked matmul (tests/test_spmm.py:9-10, tests/test_sddmm.py:8-13)"}


def reference_kernel_leg(cases, dev, torch):
    """Sputnik's own kernels (unmodified source, sm_100a build in oracle/_ref) on the same inputs."""
    try:
        from oracle import ref_gpu
        if not ref_gpu.available():
            return None
        rows, total_ms, flops = [], 0.0, 0.0
        for c in cases:
            d = c["devt"]
            nnz = c["csr"].nnz
            fns = (("spmm", lambda: ref_gpu.spmm(DIM, DIM, d["values"], d["row_indices"], d["row_offsets"],
                                                 d["column_indices"], d["dense"])),
                   ("sddmm", lambda: ref_gpu.sddmm(DIM, DIM, d["row_indices"], d["row_offsets"],
                                                   d["column_indices"], d["lhs"], d["rhs"])))
            for op, fn in fns:
                for _ in range(2):
                    fn()
                torch.cuda.synchronize()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(10):
                    fn()
                e.record()
                torch.cuda.synchronize()
                ms = s.elapsed_time(e) / 10
                rows.append({"op": op, "density": c["density"], "ms": round(ms, 4),
                             "useful_gflops": round(2.0 * nnz * DIM / (ms * 1e-3) / 1e9, 1)})
                total_ms += ms
                flops += 2.0 * nnz * DIM
        return {"what": "sputnik::CudaSpmm / CudaSddmm compiled unmodified for sm_100a (includes its "
                        "torch::zeros pre-clear as in src/spmm_cuda.cu:46)",
                "value": flops / (total_ms * 1e-3) / 1e9, "unit": UNIT, "ms_per_step": total_ms, "cases": rows}
    except Exception as exc:  # measurement aid only
        return {"unavailable": repr(exc)}


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-ref-kernel", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference_arm(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
