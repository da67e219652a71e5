#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spmm.py tests/test_gpu_modules.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_quick.log
timeout 600 python tools/bench_configs.py --config c4 > gpurun_out/configs_c4_n1.jsonl 2> gpurun_out/configs_c4_n1.err; echo "c4 rc=$?"; tail -2 gpurun_out/configs_c4_n1.err; cut -c1-200 gpurun_out/configs_c4_n1.jsonl
