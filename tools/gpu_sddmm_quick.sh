#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py -x -q > gpurun_out/pytest_sddmm.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_sddmm.log
for d in 0.05 0.1 0.15 0.25; do
  for v in 1 3; do TSP_VARIANT=$v python tools/prof_case.py sddmm $d 5; done
done
