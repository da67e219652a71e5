#!/bin/bash
set -u
python bench.py --steps 5 --no-cpu-baseline --no-ref-kernel 2>/dev/null | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print(d['value'], [ (c['op'],c['density'],c['ms']) for c in d['cases']])"
