#!/usr/bin/env python
"""bench.py -- the headline benchmark of BASELINE.json on B200.

Metric: useful GFLOP/s of CSR SpMM + SDDMM at M=N=K=4096, density 0.05/0.1/0.25/0.5, fp32
(useful FLOPs = 2*nnz*N resp. 2*nnz*K, the convention of Sputnik's own benchmark,
third_party/sputnik/sputnik/spmm/spmm_benchmark.cu.cc:29-33), next to algorithmic HBM GB/s and
the fraction of the measured HBM roofline.

A "step" = one pass of the hot path over the workload: for each of the 4 densities one SpMM and
one SDDMM through the `torch_sputnik` operator API (8 operator calls).  At N > 1 GPUs the path is
sharded with no collective -- SpMM over dense-operand columns, SDDMM over mask rows -- every rank
owning a full 4096^3 shard (weak scaling; the global problem is 4096 x 4096 x 4096*N).

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


# dram__bytes_read.sum + dram__bytes_write.sum per launch of the kernel behind (op, density), from the
# `ncu --set full` captures summarised under profiles/ (r01_*_ncu_summary.txt); None = not captured.
NCU_DRAM_TRAFFIC = {
    ("spmm", 0.5): 560.6e6,     # profiles/r01_spmm_staged_d050_ncu_summary.txt
    ("sddmm", 0.5): 1260.6e6,   # profiles/r01_sddmm_tiled2_d050_ncu_summary.txt
    ("spmm", 0.05): 112.8e6,    # profiles/r01_spmm_gather_d005_ncu_summary.txt
    ("sddmm", 0.05): 211.5e6,   # profiles/r01_sddmm_sampled_v2_d005_ncu_summary.txt
}

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
        "config": {"workload": "C2: SpMM+SDDMM M=N=K=4096 fp32 CSR, density 0.05/0.1/0.25/0.5",
                   "densities": list(DENSITIES), "impl": "reference CPU path: dense torch masked matmul"},
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
            "config": {"workload": "C2: SpMM+SDDMM M=N=K=4096 fp32 CSR, density 0.05/0.1/0.25/0.5 "
                                   "(BASELINE.json configs[1]); 8 operator calls per step",
                       "densities": list(DENSITIES),
                       "sharding": "none" if world == 1 else
                                   f"SpMM over dense columns, SDDMM over mask rows; one 4096^3 shard per rank x{world}",
                       "l2": "inputs larger than L2: 8 cases with distinct buffers (~1.7 GB) cycled per step",
                       "api": "torch_sputnik.spmm / torch_sputnik.sddmm"},
            "clocks": clocks,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d),
                    "d2h_bytes_per_step": int(d2h), "ms_per_step": e2e_ms, "steps": e2e_steps,
                    "api": "torch_sputnik_b200.host.HostPipeline (pinned host operands; upload / kernel / "
                           "download streams overlapped across calls)",
                    "serial_ms_per_step": e2e_serial_ms, "numa_bound_cores": len(numa_cores),
                    "pcie_gbs": round((h2d + d2h) / (e2e_ms * 1e-3) / 1e9, 1)},
            "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "kernel": f"{top['op']} d={top['density']}",
                         "achieved": top["algorithmic_gbs"], "peak": peak, "peak_kind": peak_kind,
                         "unit": "GB/s", "frac": top["hbm_frac"],
                         "traffic": NCU_DRAM_TRAFFIC.get((top["op"], top["density"])),
                         "algorithmic_bytes": top["algorithmic_bytes"],
                         "fp32_frac": top["fp32_frac"], "avg_launch_ms": top["ms"],
                         "note": "unstructured CSR at these shapes is bound by shared-memory/L1 gather "
                                 "bandwidth and fp32 issue, not HBM (DESIGN.md); fp32_frac is vs "
                                 f"{fp32_peak:.1f} TFLOP/s non-tensor peak"},
            "cases": table,
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
                fn()
                torch.cuda.synchronize()
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                for _ in range(3):
                    fn()
                e.record()
                torch.cuda.synchronize()
                ms = s.elapsed_time(e) / 3
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
