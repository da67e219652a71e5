#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_f16.py tests/test_gpu_band.py tests/test_gpu_attention.py -q 2>&1 | tail -4
timeout 600 python tools/ab_band.py 2>&1 | tail -1 | cut -c1-700
