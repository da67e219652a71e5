#!/usr/bin/env python
"""Secondary benchmark: the BASELINE.json configs other than the headline C2 (bench.py), one JSON
line per operator with algorithmic bytes / FLOPs, achieved GB/s and the fraction of the measured
HBM roofline.  These are the HBM-bound members of the path.

    python tools/bench_configs.py [--config c1|c3|c4|all] [--replicas R] [--gpus N]

C1  SpMM 64^3 d=0.25 and 72x64x72 sparsity 0.9 (launch-latency bound)
C3  sparse_linear fwd + bwd: W 4096x4096 at 90 % sparsity, 8192 tokens as [16, 512, 4096]
C4  sparse attention: R = batch*heads replicas (default 128; sharded over --gpus), seq 4096,
    head_dim 64, mask 90 % sparse: SDDMM -> scaled CSR softmax -> SpMM (+ csr_transpose, permute)
Under torchrun (WORLD_SIZE > 1) C4 is sharded over replicas and C3 over tokens with one NCCL
all-reduce of values.grad; rank 0 prints max-over-ranks times.
"""
from __future__ import annotations

import argparse
import time
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from oracle.csr import dense_operand, random_csr  # noqa: E402  (input generator only)
from torch_sputnik_b200 import dist as tdist  # noqa: E402

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


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]), "measured"
    except Exception:
        return 6650.0, "fallback"


WORLD = int(os.environ.get("WORLD_SIZE", "1"))
RANK = int(os.environ.get("RANK", "0"))
LOCAL = int(os.environ.get("LOCAL_RANK", "0"))


def _loop_ms(body, reps):
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    s.record()
    for _ in range(reps):
        body()
    e.record()
    e.synchronize()
    return s.elapsed_time(e) / reps


def timeit(fn, reps=20, warm=5, flush=None):
    """Mean device time of fn() in ms over `reps` back-to-back launches (CUDA events on the launching
    stream, no host sync inside the loop, so host launch latency is pipelined as in real use).
    `flush` (a tensor larger than L2) is rewritten between calls when the working set is smaller
    than L2; its own time, measured the same way, is subtracted."""
    for _ in range(warm):
        fn()
        if flush is not None:
            flush.add_(1.0)   # first use loads the elementwise kernel's module: keep that out of the timed loop
    # the host-side setup between measurements (numpy generators, reference builds) lets the SM clock
    # drop; keep the GPU busy for >= 150 ms before the timed loop so that it runs at the boost clock
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    while time.perf_counter() - t0 < 0.15:
        for _ in range(warm):
            fn()
        torch.cuda.synchronize()
    if flush is None:
        ms = _loop_ms(fn, reps)
    else:
        both = _loop_ms(lambda: (flush.add_(1.0), fn()), reps)
        ms = both - _loop_ms(lambda: flush.add_(1.0), reps)
    if WORLD > 1:
        t = torch.tensor([ms], device="cuda", dtype=torch.float64)
        torch.distributed.all_reduce(t, op=torch.distributed.ReduceOp.MAX)
        ms = float(t)
    return ms


def emit(config, op, ms, nbytes, flops, extra=None):
    if RANK != 0:
        return
    peak, kind = peak_gbs()
    gbs = nbytes * WORLD / (ms * 1e-3) / 1e9
    line = {"config": config, "op": op, "ms": round(ms, 4), "n_gpus": WORLD,
            "algorithmic_bytes": int(nbytes * WORLD), "algorithmic_gbs": round(gbs, 1),
            "hbm_frac_per_gpu": round(gbs / WORLD / peak, 4), "peak_gbs": peak, "peak_kind": kind,
            "useful_gflops": round(flops * WORLD / (ms * 1e-3) / 1e9, 1)}
    if extra:
        line.update(extra)
    emit_line(json.dumps(line))


def ref_ms(fn, reps=5):
    """Time the reference's own kernel (oracle/_ref, measurement aid) when the library travelled."""
    try:
        from oracle import ref_gpu
        if not ref_gpu.available() or WORLD > 1:
            return None
        fn(ref_gpu)
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(reps):
            fn(ref_gpu)
        e.record()
        e.synchronize()
        return round(s.elapsed_time(e) / reps, 4)
    except Exception as exc:
        return repr(exc)


def dev_csr(a, dev):
    t = lambda x: torch.from_numpy(np.ascontiguousarray(x)).to(dev)
    return t(a.values), t(a.row_indices), t(a.row_offsets), t(a.column_indices)


def c1(dev, flush):
    for name, (m, k, n, d) in {"64^3 d=0.25": (64, 64, 64, 0.25), "72x64x72 d=0.1": (72, 64, 72, 0.1)}.items():
        a = random_csr(m, k, d, seed=0, values="u01", round_to=4)
        v, ri, ro, ci = dev_csr(a, dev)
        b = torch.from_numpy(dense_operand((k, n), 1)).to(dev)
        ms = timeit(lambda: ops.spmm(m, k, v, ri, ro, ci, b), reps=200, warm=20)
        emit("C1", f"spmm {name}", ms, 8 * a.nnz + 4 * (2 * m + 1) + 4 * k * n + 4 * m * n, 2.0 * a.nnz * n,
             {"note": "launch-latency bound; time includes the torch shim"})


def c3(dev, flush):
    m = k = 4096
    batch, tokens = 16 // WORLD if 16 % WORLD == 0 else 16, 512
    batch = int(os.environ.get("TSP_C3_BATCH", batch))
    a = random_csr(m, k, 0.1, seed=0, values="pm1")
    v, ri, ro, ci = dev_csr(a, dev)
    nnz = a.nnz
    x = torch.from_numpy(dense_operand((batch, k, tokens), 1 + RANK, "pm1")).to(dev)      # x^T: [B, in, S]
    dy = torch.from_numpy(dense_operand((batch, m, tokens), 2 + RANK, "pm1")).to(dev)     # [B, out, S]
    n_tot = batch * tokens
    # forward: left_spmm
    ms = timeit(lambda: ops.left_spmm(m, k, v, ri, ro, ci, x), flush=flush)
    emit("C3", "fwd left_spmm", ms, 8 * nnz + 4 * (2 * m + 1) + 4 * k * n_tot + 4 * m * n_tot, 2.0 * nnz * n_tot,
         {"ref_sputnik_ms": ref_ms(lambda r: r.spmm(m, k, v, ri, ro, ci, x, values_shared=True))})
    # weight gradient: replicated sddmm over the batch, then the sum over replicas
    ms = timeit(lambda: ops.sddmm_reduce(m, k, ri, ro, ci, dy, x), flush=flush)
    emit("C3", "wgrad sddmm_reduce", ms, 4 * nnz + 4 * (2 * m + 1) + 4 * m * n_tot + 4 * k * n_tot + 4 * nnz, 2.0 * nnz * n_tot,
         {"ref_sputnik_ms": ref_ms(lambda r: r.sddmm(m, k, ri, ro, ci, dy, x).sum(0))})
    # transpose of the static topology (once) and the per-step value permutation
    ms = timeit(lambda: ops.csr_transpose(m, k, v, ro, ci), flush=flush)
    emit("C3", "csr_transpose", ms, 16 * nnz + 4 * (m + 1) + 4 * (k + 1), 0.0,
         {"ref_cusparse_ms": ref_ms(lambda r: r.csr_transpose(m, k, v, ro, ci))})
    vt, rot, cit, perm = ops.csr_transpose_with_perm(m, k, v, ro, ci)
    ms = timeit(lambda: ops.permute_values(v, perm), flush=flush)
    emit("C3", "permute_values (cached transpose)", ms, 12 * nnz, 0.0)
    rit = torch.argsort(rot[1:] - rot[:-1], descending=True, stable=True).to(torch.int32)
    ms = timeit(lambda: ops.left_spmm(k, m, vt, rit, rot, cit, dy), flush=flush)
    emit("C3", "dgrad left_spmm", ms, 8 * nnz + 4 * (2 * k + 1) + 4 * m * n_tot + 4 * k * n_tot, 2.0 * nnz * n_tot)
    if os.environ.get("TSP_C3_REPEAT"):
        for tag, rows in (("swizzled", ri), ("identity", torch.arange(m, dtype=torch.int32, device=dev))):
            ms = timeit(lambda: ops.left_spmm(m, k, v, rows, ro, ci, x), flush=flush)
            emit("C3", f"fwd left_spmm (repeat, {tag} rows)", ms, 0, 2.0 * nnz * n_tot)
    if WORLD > 1:
        p = torch.nn.Parameter(v.clone())
        bucket = tdist.ValueGradBucket([p])
        ms = timeit(lambda: bucket.all_reduce().wait())
        emit("C3", "allreduce values.grad (NCCL)", ms, 4 * nnz, 0.0, {"bytes_per_rank": 4 * nnz})


def c4(dev, flush, replicas):
    s, d = 4096, 64
    reps = tdist.shard_replicas(replicas, WORLD, RANK)
    r = reps.stop - reps.start
    a = random_csr(s, s, 0.1, seed=0, values="pm1", round_to=4)   # nnz % 4 == 0: the reference kernel needs it for 3-D values
    _, ri, ro, ci = dev_csr(a, dev)
    nnz = a.nnz
    g = torch.Generator(device="cpu").manual_seed(10 + RANK)
    q, kk, vv = (torch.rand((r, s, d), generator=g).mul_(2).sub_(1).to(dev) for _ in range(3))
    topo = 4 * nnz + 4 * (2 * s + 1)
    scores = ops.sddmm(s, s, ri, ro, ci, q, kk)
    ms_sddmm = timeit(lambda: ops.sddmm(s, s, ri, ro, ci, q, kk))
    emit("C4", "sddmm", ms_sddmm, r * (4 * nnz + 8 * s * d) + topo, 2.0 * nnz * d * r,
         {"replicas_per_gpu": r, "ref_sputnik_ms": ref_ms(lambda g: g.sddmm(s, s, ri, ro, ci, q, kk), 2)})
    probs = ops.sparse_softmax_scaled(scores, ri, ro, ci, 0.125)
    ms_soft = timeit(lambda: ops.sparse_softmax_scaled(scores, ri, ro, ci, 0.125))
    emit("C4", "csr_softmax (scale fused)", ms_soft, r * 8 * nnz + 4 * (2 * s + 1), 5.0 * nnz * r,
         {"replicas_per_gpu": r, "ref_sputnik_ms": ref_ms(lambda g: g.sparse_softmax(scores, ri, ro, ci), 2)})
    ms_spmm = timeit(lambda: ops.spmm(s, s, probs, ri, ro, ci, vv))
    emit("C4", "spmm", ms_spmm, r * (4 * nnz + 8 * s * d) + topo, 2.0 * nnz * d * r,
         {"replicas_per_gpu": r, "ref_sputnik_ms": ref_ms(lambda g: g.spmm(s, s, probs, ri, ro, ci, vv), 2)})
    dy = torch.rand_like(probs)
    ms_bwd = timeit(lambda: ops.sparse_softmax_backward(probs, dy, ri, ro, ci, 0.125))
    emit("C4", "csr_softmax backward", ms_bwd, r * 12 * nnz + 4 * (2 * s + 1), 4.0 * nnz * r, {"replicas_per_gpu": r})
    _, rot, cit, perm = ops.csr_transpose_with_perm(s, s, probs[0].contiguous(), ro, ci)
    ms_perm = timeit(lambda: ops.permute_values(probs, perm))
    emit("C4", "permute_values [R,nnz]", ms_perm, r * 8 * nnz + 4 * nnz, 0.0, {"replicas_per_gpu": r})
    rit = torch.argsort(rot[1:] - rot[:-1], descending=True, stable=True).to(torch.int32)
    ms_t2 = timeit(lambda: ops.spmm(s, s, ops.permute_values(probs, perm), rit, rot, cit, vv))
    ms_t1 = timeit(lambda: ops.spmm_perm(s, s, probs, perm, rit, rot, cit, vv))
    emit("C4", "A^T spmm (backward dK/dV): spmm_perm, values read through the cached permutation", ms_t1,
         r * (4 * nnz + 8 * s * d) + 2 * topo, 2.0 * nnz * d * r,
         {"replicas_per_gpu": r, "permute_values_then_spmm_ms": round(ms_t2, 4)})
    total = ms_sddmm + ms_soft + ms_spmm
    emit("C4", "attention fwd (3 ops)", total, r * (16 * nnz + 16 * s * d) + 2 * topo + 4 * (2 * s + 1),
         4.0 * nnz * d * r, {"replicas_per_gpu": r, "replicas_total": replicas})
    # A/B: the L1-gather kernels the band kernels replace at this shape
    ops.set_variant("spmm", 1)
    ms_g = timeit(lambda: ops.spmm(s, s, probs, ri, ro, ci, vv))
    ops.set_variant("spmm", 0)
    ops.set_variant("sddmm", 5)
    ms_r = timeit(lambda: ops.sddmm(s, s, ri, ro, ci, q, kk))
    ops.set_variant("sddmm", 0)
    emit("C4", "spmm (gather kernel, fp32)", ms_g, r * (4 * nnz + 8 * s * d) + topo, 2.0 * nnz * d * r,
         {"replicas_per_gpu": r, "band_ms": round(ms_spmm, 4)})
    emit("C4", "sddmm (rows kernel, fp32)", ms_r, r * (4 * nnz + 8 * s * d) + topo, 2.0 * nnz * d * r,
         {"replicas_per_gpu": r, "band_ms": round(ms_sddmm, 4)})
    # 16-bit storage, fp32 arithmetic: the same three ops (algorithmic bytes halve, topology stays int32)
    for name, dt in (("f16", torch.float16), ("bf16", torch.bfloat16)):
        qh, kh, vh = q.to(dt), kk.to(dt), vv.to(dt)
        sc_h = ops.sddmm(s, s, ri, ro, ci, qh, kh)
        ms_a = timeit(lambda: ops.sddmm(s, s, ri, ro, ci, qh, kh))
        pr_h = ops.sparse_softmax_scaled(sc_h, ri, ro, ci, 0.125)
        ms_b = timeit(lambda: ops.sparse_softmax_scaled(sc_h, ri, ro, ci, 0.125))
        ms_c = timeit(lambda: ops.spmm(s, s, pr_h, ri, ro, ci, vh))
        emit("C4", f"sddmm {name}", ms_a, r * (2 * nnz + 4 * s * d) + topo, 2.0 * nnz * d * r, {"replicas_per_gpu": r})
        emit("C4", f"csr_softmax {name} (scale fused)", ms_b, r * 4 * nnz + 4 * (2 * s + 1), 5.0 * nnz * r, {"replicas_per_gpu": r})
        emit("C4", f"spmm {name}", ms_c, r * (2 * nnz + 4 * s * d) + topo, 2.0 * nnz * d * r, {"replicas_per_gpu": r})
        emit("C4", f"attention fwd (3 ops) {name}", ms_a + ms_b + ms_c, r * (8 * nnz + 8 * s * d) + 2 * topo + 4 * (2 * s + 1),
             4.0 * nnz * d * r, {"replicas_per_gpu": r, "replicas_total": replicas, "fp32_chain_ms": round(total, 4)})
        del qh, kh, vh, sc_h, pr_h
    for w in [int(x) for x in os.environ.get("TSP_ATTN_WARPS", "0").split(",")]:
        ops.set_variant("attn_warps", w)
        ms_fused = timeit(lambda: ops.sparse_attention(s, s, ri, ro, ci, q, kk, vv, 0.125))
        emit("C4", "attention fwd fused (1 kernel)" + (f" warps={w}" if w else ""), ms_fused,
             r * 16 * s * d + topo + 4 * r * s, 4.0 * nnz * d * r,
             {"replicas_per_gpu": r, "replicas_total": replicas, "three_op_chain_ms": round(total, 4)})
    ops.set_variant("attn_warps", 0)


def main():
    protect_stdout()
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="all")
    ap.add_argument("--replicas", type=int, default=128)
    ap.add_argument("--gpus", type=int, default=1)
    args = ap.parse_args()
    torch.cuda.set_device(LOCAL)
    dev = torch.device("cuda", LOCAL)
    if WORLD > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL's version banner must not land on stdout
        torch.distributed.init_process_group("nccl", device_id=dev)
    flush = torch.zeros(64 * 1024 * 1024, device=dev)   # 256 MB > L2
    if args.config in ("c1", "all"):
        c1(dev, flush)
    if args.config in ("c3", "all"):
        c3(dev, flush)
    if args.config in ("c4", "all"):
        c4(dev, flush, args.replicas)
    if WORLD > 1:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
