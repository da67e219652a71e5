#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_softmax_transpose.py tests/test_gpu_plans.py tests/test_gpu_many_mask.py -q -x 2>&1 | tail -5
timeout 300 python tools/ab_transpose.py 2>&1 | tee gpurun_out/r02_transpose_ab.txt
