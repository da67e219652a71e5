#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_sddmm.py -q -x 2>&1 | tail -2
python tools/prof_case.py sddmm 0.5 5 | tail -1
python tools/prof_case.py sddmm 0.25 5 | tail -1
python tools/prof_case.py spmm 0.5 2 | tail -1
rm -f gpurun_out/tiled2_src.ncu-rep
timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -c 1 -f -o gpurun_out/staged_src python tools/prof_case.py spmm 0.5 2 > gpurun_out/ncu_staged_src.log 2>&1; echo "rc=$?"
