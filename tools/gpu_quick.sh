#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_sddmm.py tests/test_gpu_f16.py -q -x 2>&1 | tail -2
for d in 0.02 0.05 0.1; do
  python tools/prof_case.py sddmm $d 5 | tail -1
  TSP_DTYPE=f16 python tools/prof_case.py sddmm $d 5 | tail -1
done
python tools/bench_configs.py --config C3 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    x=json.loads(l); print(x.get('op'), x.get('ms'))"
