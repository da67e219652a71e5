#!/bin/bash
# scratch: the command of the moment for a single `gpurun -- 'bash tools/gpu_quick.sh'`
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q --deselect tests/test_gpu_ref_callers.py > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 900 python -m pytest tests/test_gpu_ref_callers.py -q > gpurun_out/pytest_ref_callers.log 2>&1; echo "ref callers rc=$?" | tee -a gpurun_out/pytest_ref_callers.log
tail -40 gpurun_out/pytest_ref_callers.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench.json 2> gpurun_out/bench.err; rc=$?; echo "bench rc=$rc"
tail -c 1500 gpurun_out/bench.err; cat gpurun_out/bench.json
