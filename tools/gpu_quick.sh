#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py -x -q -k rows > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_quick.log
for v in 5 6; do for w in 0 16 8; do echo "variant $v warps $w"; TSP_VARIANT=$v TSP_SDDMM_WARPS=$w python tools/prof_c4.py sddmm 128 4; done; done
