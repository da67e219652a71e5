#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_modules.py -x -q > gpurun_out/pytest_quick.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_quick.log
python tools/train_dp_transformer.py --steps 10 > gpurun_out/c5_n1.json 2> gpurun_out/c5_n1.err; echo rc=$?; tail -5 gpurun_out/c5_n1.err; cat gpurun_out/c5_n1.json
