#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_plans.py tests/test_gpu_modules.py tests/test_gpu_ref_callers.py -q 2>&1 | tail -5
timeout 300 python tools/train_dp_transformer.py --graph 1 --steps 5 2>&1 | tail -1 | cut -c1-200
