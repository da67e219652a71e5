#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_sddmm.py tests/test_gpu_spmm.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_quick.log
for w in 4 16 32; do echo warps=$w; TSP_SPMM_WARPS=$w python tools/prof_c4.py spmm 16 3; TSP_SDDMM_WARPS=$w python tools/prof_c4.py sddmm 16 3; done
python tools/bench_configs.py --config c4 > gpurun_out/configs_c4.jsonl 2> gpurun_out/configs_c4.err; echo rc=$?; tail -3 gpurun_out/configs_c4.err
