#!/usr/bin/env python
"""BASELINE.json configs[4]: data-parallel training step of the sparse Transformer
(torch_sputnik_b200/modules/transformer.py) -- 90 %-sparse linear layers, shared-mask sparse
attention -- with ONE NCCL all-reduce of the nnz-long value gradients per step (plus one for the
small dense parameters).  One process per GPU:

    python tools/train_dp_transformer.py [--steps 10] [--batch 4]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \\
        --master-port 29500 tools/train_dp_transformer.py

Prints one JSON line on rank 0: tokens/s over all ranks (weak scaling: --batch sequences per GPU),
ms per step (max over ranks, CUDA events), bytes all-reduced per step.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch_sputnik_b200  # noqa: E402,F401
import torch_sputnik as ops  # noqa: E402
from torch_sputnik_b200 import dist as tdist  # noqa: E402
from torch_sputnik_b200.modules.transformer import SparseTransformer  # noqa: E402


def main():
    real_out = os.fdopen(os.dup(1), "w")   # fd 1 -> stderr from here on: NCCL prints its banner to fd 1
    sys.stdout.flush()
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--layers", type=int, default=6)
    ap.add_argument("--hidden", type=int, default=512)
    ap.add_argument("--heads", type=int, default=8)
    ap.add_argument("--ffn", type=int, default=2048)
    ap.add_argument("--seq", type=int, default=512)
    ap.add_argument("--lr", type=float, default=1e-2)
    ap.add_argument("--graph", type=int, default=1,
                    help="1: capture the whole training step (forward, backward, both all-reduces, update) in ONE "
                         "CUDA graph and replay it; 0: eager launches")
    args = ap.parse_args()

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # NCCL's version banner must not land on stdout
        torch.distributed.init_process_group("nccl", device_id=dev)

    torch.manual_seed(0)                       # identical initial weights and masks on every rank
    vocab = 1000
    model = SparseTransformer(args.layers, args.hidden, args.heads, args.ffn, args.seq, vocab, device=dev).to(dev)
    model.setup_sparse()
    sparse = tdist.sparse_value_parameters(model)
    sparse_ids = {id(p) for p in sparse}
    dense = [p for p in model.parameters() if p.requires_grad and id(p) not in sparse_ids]
    tdist.broadcast_parameters(sparse + dense)
    b_sparse = tdist.ValueGradBucket(sparse)
    b_dense = tdist.ValueGradBucket(dense)

    g = torch.Generator(device="cpu").manual_seed(100 + rank)   # each rank its own token shard
    tokens = torch.randint(0, vocab, (args.batch, args.seq), generator=g).to(dev)
    target = torch.randint(0, vocab, (args.batch, args.seq), generator=g).to(dev)
    loss_fn = torch.nn.CrossEntropyLoss()

    def step():
        b_sparse.zero_()
        b_dense.zero_()
        logits = model(tokens)
        loss = loss_fn(logits.reshape(-1, vocab), target.reshape(-1))
        loss.backward()
        b_sparse.all_reduce()                  # side stream: overlaps the dense bucket's launch
        b_dense.all_reduce()
        b_sparse.wait()
        b_dense.wait()
        with torch.no_grad():
            for p in sparse + dense:
                p.add_(p.grad, alpha=-args.lr)
        return loss

    losses = []
    for _ in range(args.warmup):
        losses.append(float(step()))
    eager_step = step
    l_before = ops.launch_count()
    eager_step()
    launches_per_step = ops.launch_count() - l_before
    if args.graph:
        # The step is ~1000 short launches (348 of ours + autograd's elementwise / layer-norm / embedding kernels) and is
        # bound by host launch latency, not by the GPU.  Capture it once: every buffer it touches is static (tokens,
        # target, parameters, the two gradient buckets; temporaries come from the graph's private pool; the transposed
        # topologies were cached by the warm-up steps), the NCCL all-reduces on the buckets' side streams are captured
        # as part of the same graph (fork / join through wait_stream).
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(2):
                eager_step()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            static_loss = eager_step()

        def step():
            graph.replay()
            return static_loss
        losses.append(float(step()))
    if world > 1:
        torch.distributed.barrier()
    torch.cuda.synchronize()
    l0 = ops.launch_count()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(args.steps):
        loss = step()
    e.record()
    torch.cuda.synchronize()
    losses.append(float(loss))
    ms = torch.tensor([s.elapsed_time(e) / args.steps], device=dev, dtype=torch.float64)
    if world > 1:
        torch.distributed.all_reduce(ms, op=torch.distributed.ReduceOp.MAX)
    if rank == 0:
        ms = float(ms)
        real_out.write(json.dumps({
            "config": "C5: data-parallel sparse Transformer, 90% sparse linear layers + shared-mask sparse attention",
            "n_gpus": world, "ms_per_step": round(ms, 3), "cuda_graph": bool(args.graph),
            "tokens_per_s": round(world * args.batch * args.seq / (ms * 1e-3), 1), "scaling": "weak",
            "model": {"layers": args.layers, "hidden": args.hidden, "heads": args.heads, "ffn": args.ffn,
                      "seq": args.seq, "batch_per_gpu": args.batch},
            "allreduce_bytes_per_step": {"sparse_values": b_sparse.nbytes, "dense_params": b_dense.nbytes,
                                         "dense_equivalent_of_sparse": 10 * b_sparse.nbytes},
            "tsp_kernel_launches_per_step": launches_per_step,
            "loss_first_last": [round(losses[0], 4), round(losses[-1], 4)]}) + "\n")
        real_out.flush()
    if world > 1:
        torch.cuda.synchronize()
        if args.graph:
            # A captured graph holds NCCL work: release it before the communicator goes away, and leave without the
            # interpreter's teardown -- with the graph alive, barrier + destroy_process_group did not return on 8 GPUs
            # (the result line above had been written; the run sat in teardown until the job limit).
            graph.reset()
            del graph
            sys.stderr.flush()
            os._exit(0)
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
