#!/bin/bash
set -u
mkdir -p gpurun_out
for dt in f16 f32; do
TSP_DTYPE=$dt timeout 300 ncu --set full --clock-control none --profile-from-start off -c 1 -f -o gpurun_out/r02_softmax_$dt python tools/prof_c4.py softmax 128 2 > /dev/null 2>&1
python tools/ncu_summary.py gpurun_out/r02_softmax_$dt.ncu-rep | grep -E "==|time_duration|dram__bytes|dram__thr|warps_active|lsu_wavefronts.avg|issue_active|inst_executed.sum|scoreboard|lg_throttle|occupancy_limit|registers|lts__t_sector_hit|mio_thr"
done
rm -f gpurun_out/r02_softmax_*.ncu-rep
