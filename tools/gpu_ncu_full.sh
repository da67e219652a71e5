#!/bin/bash
# ncu --set full captures of the dominant kernels (one op per capture; the plain command runs first).
set -u
mkdir -p gpurun_out
cap() {  # name op density kernel-regex
  python tools/prof_case.py $2 $3 3 > gpurun_out/plain_$1.log 2>&1 || { echo "plain $1 failed"; return; }
  cat gpurun_out/plain_$1.log
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 1 -c 1 -f -o gpurun_out/$1 \
    python tools/prof_case.py $2 $3 3 > gpurun_out/ncu_$1.log 2>&1; echo "ncu $1 rc=$?"
}
cap r01_sddmm_tiled_d050 sddmm 0.5 sddmm_tiled_kernel
cap r01_spmm_staged_d050 spmm 0.5 spmm_staged_kernel
cap r01_sddmm_gather_d005 sddmm 0.05 sddmm_gather_kernel
cap r01_spmm_gather_d005 spmm 0.05 spmm_gather_kernel
ls -la gpurun_out
