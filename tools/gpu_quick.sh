#!/bin/bash
# scratch: the command of the moment for a single `gpurun -- 'bash tools/gpu_quick.sh'`
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
