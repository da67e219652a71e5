#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_plans.py tests/test_gpu_softmax_transpose.py tests/test_gpu_many_mask.py tests/test_gpu_ref_callers.py tests/test_gpu_modules.py -q -x 2>&1 | tail -8
timeout 300 python tools/ab_permute.py 2>&1 | tee gpurun_out/r02_permute_ab.txt
