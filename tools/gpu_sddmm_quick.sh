#!/bin/bash
set -u
timeout 900 python -m pytest tests/test_gpu_sddmm.py -x -q 2>&1 | tail -2
for d in 0.25 0.5; do TSP_VARIANT=2 python tools/prof_case.py sddmm $d 4; done
