#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_f16.py tests/test_gpu_sddmm.py tests/test_gpu_plans.py tests/test_gpu_sharding.py -q -x 2>&1 | tail -8
timeout 600 python tools/ab_sddmm_f16.py 2>&1 | tee gpurun_out/r02_sddmm_f16_ab.txt
