#!/bin/bash
set -u
for n in 512 1024 4096; do python tools/prof_case.py spmm 0.25 4 $n | tail -1; done
timeout 900 python -m pytest tests/test_gpu_many_mask.py tests/test_gpu_ref_callers.py tests/test_gpu_spmm.py tests/test_gpu_modules.py tests/test_gpu_sharding.py -q 2>&1 | tail -4
