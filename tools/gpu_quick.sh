#!/bin/bash
# Scratch script for one short gpurun call: the parity tests of the kernels being worked on + a timing line.
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py tests/test_gpu_spmm.py -q -x 2>&1 | tail -2
for d in 0.05 0.5; do python tools/prof_case.py sddmm $d 5 | tail -1; python tools/prof_case.py spmm $d 5 | tail -1; done
