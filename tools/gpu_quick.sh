#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py -q -x -k tiled3 2>&1 | tail -2
python tools/prof_case.py sddmm 0.5 3 | tail -1
for v in 1 2; do TSP_SDDMM_TILED3=$v python tools/prof_case.py sddmm 0.5 3 | tail -1; done
