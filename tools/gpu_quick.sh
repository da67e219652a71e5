#!/bin/bash
set -u
mkdir -p gpurun_out
TSP_C3_BATCH=2 TSP_C3_REPEAT=1 timeout 600 python tools/bench_configs.py --config c3 2>&1 | cut -c1-100
