#!/bin/bash
# ncu --set full capture of ONE steady-state launch of every kernel bench.py dispatches (C2: 8 cases) and of
# the C4 attention kernels; summaries + the traffic table are written next to the reports in gpurun_out/.
# Each ncu pass runs only after the same command has exited 0 without ncu.
set -u
mkdir -p gpurun_out
cap() {  # name  command...
  local name=$1; shift
  "$@" > gpurun_out/plain_$name.log 2>&1 || { echo "plain $name failed"; cat gpurun_out/plain_$name.log; return; }
  tail -1 gpurun_out/plain_$name.log
  timeout 600 ncu --set full --clock-control none --import-source on --profile-from-start off -c 1 -f \
    -o gpurun_out/$name "$@" > gpurun_out/ncu_$name.log 2>&1; echo "ncu $name rc=$?"
}
for d in 0.05 0.1 0.25 0.5; do
  tag=$(python -c "print('%03d' % round($d*1000))")
  cap r02_c2_spmm_d$tag python tools/prof_case.py spmm $d 2
  cap r02_c2_sddmm_d$tag python tools/prof_case.py sddmm $d 2
done
TSP_DTYPE=f16 cap r02_c2_sddmm_f16_d100 python tools/prof_case.py sddmm 0.1 2
TSP_DTYPE=f16 cap r02_c2_spmm_f16_d100 python tools/prof_case.py spmm 0.1 2
cap r02_c4_sddmm_band python tools/prof_c4.py sddmm 32 2
TSP_VARIANT=5 cap r02_c4_sddmm_rows python tools/prof_c4.py sddmm 32 2
cap r02_c4_spmm_gather python tools/prof_c4.py spmm 32 2
cap r02_c4_softmax python tools/prof_c4.py softmax 32 2
TSP_DTYPE=f16 cap r02_c4_sddmm_rows_f16 python tools/prof_c4.py sddmm 32 2
TSP_DTYPE=f16 cap r02_c4_spmm_rowsplit_f16 python tools/prof_c4.py spmm 32 2
python tools/ncu_traffic.py gpurun_out gpurun_out > gpurun_out/traffic.log 2>&1; tail -3 gpurun_out/traffic.log
rm -f gpurun_out/r02_c2_*.ncu-rep gpurun_out/r02_c4_spmm*.ncu-rep gpurun_out/r02_c4_softmax.ncu-rep gpurun_out/r02_c4_sddmm_rows*.ncu-rep
