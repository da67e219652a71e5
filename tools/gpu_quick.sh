#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py tests/test_gpu_modules.py tests/test_gpu_spmm.py -x -q 2>&1 | tail -2
python tools/prof_small_wgrad.py
for g in 0 1; do
timeout 300 python tools/train_dp_transformer.py --steps 20 --graph $g 2> gpurun_out/c5_n1_g$g.err | cut -c100-260
done
