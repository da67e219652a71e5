#!/bin/bash
set -u
mkdir -p gpurun_out
for op in spmm sddmm; do
TSP_DTYPE=f16 timeout 300 ncu --set full --clock-control none --profile-from-start off -c 1 -f -o gpurun_out/r02_c4_${op}_f16_mixed python tools/prof_c4.py $op 32 2 > /dev/null 2>&1
python tools/ncu_summary.py gpurun_out/r02_c4_${op}_f16_mixed.ncu-rep > gpurun_out/r02_c4_${op}_f16_mixed_ncu_summary.txt
grep -E "==|time_duration|registers_per|warps_active|lsu_wavefronts.avg|issue_active|pipe_fma_cycles|inst_executed.sum|long_scoreboard|short_scoreboard|wait_per|not_selected|l1tex__t_sector_hit|lts__t_sector|xbar2l1tex|lg_throttle|mio_throttle" gpurun_out/r02_c4_${op}_f16_mixed_ncu_summary.txt
done
rm -f gpurun_out/*.ncu-rep
