#!/bin/bash
# usage: gpu_ncu_one.sh NAME OP DENSITY KERNEL_REGEX [VARIANT]
set -u
mkdir -p gpurun_out
export TSP_VARIANT=${5:-}
python tools/prof_case.py $2 $3 3 > gpurun_out/plain_$1.log 2>&1 || { echo "plain failed"; cat gpurun_out/plain_$1.log; exit 1; }
cat gpurun_out/plain_$1.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 1 -c 1 -f -o gpurun_out/$1 \
  python tools/prof_case.py $2 $3 3 > gpurun_out/ncu_$1.log 2>&1; echo "ncu rc=$?"
