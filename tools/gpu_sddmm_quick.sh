#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py -x -q > gpurun_out/pytest_sddmm.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_sddmm.log
export TSP_VARIANT=3
for d in 0.02 0.05 0.1 0.15 0.2; do python tools/prof_case.py sddmm $d 4; done
