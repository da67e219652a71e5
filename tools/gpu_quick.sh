#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_spmm.py tests/test_gpu_many_mask.py tests/test_gpu_golden.py -q 2>&1 | tail -3
for d in 0.05 0.1; do python tools/prof_case.py spmm $d 5 | tail -1; done
python tools/prof_c4.py spmm 128 4 | tail -1
