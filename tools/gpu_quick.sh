#!/bin/bash
set -u
mkdir -p gpurun_out
for d in 0.25 0.5; do python tools/prof_case.py spmm $d 5 | tail -1; done
python tools/prof_case.py sddmm 0.5 5 | tail -1
