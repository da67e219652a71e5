#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_modules.py tests/test_gpu_many_mask.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_quick.log
