#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_modules.py -q 2>&1 | tail -3
timeout 600 python tools/train_dp_transformer.py --graph 0 --steps 5 2>/dev/null | cut -c1-260
timeout 600 python tools/train_dp_transformer.py --graph 1 --steps 10 2>/dev/null | cut -c1-260
