#!/bin/bash
set -u
mkdir -p gpurun_out
cap() { local name=$1; shift; "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "plain $name failed"; return; }; tail -1 gpurun_out/plain_$name.log
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -c 1 -f -o gpurun_out/$name "$@" > gpurun_out/ncu_$name.log 2>&1; echo "ncu $name rc=$?"; }
cap r02_c4_softmax python tools/prof_c4.py softmax 128 2
TSP_DTYPE=f16 cap r02_c4_softmax_f16 python tools/prof_c4.py softmax 128 2
python tools/ncu_traffic.py gpurun_out gpurun_out > gpurun_out/traffic.log 2>&1
rm -f gpurun_out/*.ncu-rep
