#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_softmax_transpose.py tests/test_gpu_f16.py tests/test_gpu_many_mask.py tests/test_gpu_attention.py -q -x 2>&1 | tail -2
python tools/prof_c4.py softmax 128 5 | tail -1
TSP_DTYPE=f16 python tools/prof_c4.py softmax 128 5 | tail -1
