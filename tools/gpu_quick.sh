#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_sddmm.py tests/test_gpu_modules.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_quick.log
python tools/bench_configs.py --config c3 > gpurun_out/configs_c3.jsonl 2> gpurun_out/configs_c3.err; echo rc=$?; tail -3 gpurun_out/configs_c3.err
