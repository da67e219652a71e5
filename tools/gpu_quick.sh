#!/bin/bash
set -u
mkdir -p gpurun_out
CUDA_LAUNCH_BLOCKING=1 python tests/golden/make_golden_gpu.py --half gpurun_out/ref_b200_f16.npz 2>&1 | tail -25
