#!/bin/bash
# usage: gpu_multi.sh N   -- the sharded paths on N GPUs of one box (run under `gpurun --gpus N`)
# Every command has a SHORT timeout: an 8-GPU call is charged 8x, and one process that does not exit (a hang in NCCL
# teardown once cost 74 GPU-minutes) must not hold the box until the call limit.
set -u
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511"
nvidia-smi -L > gpurun_out/smi_n$N.txt
timeout 240 $TR bench.py --gpus $N --steps 5 --warmup 3 --no-cpu-baseline --no-ref-kernel > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_n$N.err
timeout 240 $TR tools/bench_configs.py --config c4 > gpurun_out/configs_c4_n$N.jsonl 2> gpurun_out/configs_c4_n$N.err; echo "c4 rc=$?"; tail -2 gpurun_out/configs_c4_n$N.err
timeout 240 $TR tools/bench_configs.py --config c3 > gpurun_out/configs_c3_n$N.jsonl 2> gpurun_out/configs_c3_n$N.err; echo "c3 rc=$?"; tail -2 gpurun_out/configs_c3_n$N.err
timeout 120 $TR tools/train_dp_transformer.py --steps 10 > gpurun_out/c5_n$N.json 2> gpurun_out/c5_n$N.err; echo "c5 rc=$?"; tail -2 gpurun_out/c5_n$N.err
cat gpurun_out/c5_n$N.json
