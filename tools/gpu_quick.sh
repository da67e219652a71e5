#!/bin/bash
set -u
for d in 0.05 0.1 0.2; do
  echo "== density $d default"; python tools/prof_case.py sddmm $d 5 | tail -1
  for p in 48 40; do
    echo "-- p=$p auto bn"; TSP_SDDMM_P=$p python tools/prof_case.py sddmm $d 5 | tail -1
    for bn in 64 96 112; do echo "-- p=$p bn=$bn"; TSP_SDDMM_P=$p TSP_SDDMM_BN=$bn python tools/prof_case.py sddmm $d 5 | tail -1; done
  done
done
