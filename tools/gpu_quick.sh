#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_f16.py tests/test_gpu_band.py -q 2>&1 | tail -4
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-ref-kernel > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench.json'))
print(d['value'], d['ms_per_step'])
for c in d['cases_f16']: print(c)
PY
