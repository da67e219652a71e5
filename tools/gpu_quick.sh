#!/bin/bash
set -u
mkdir -p gpurun_out
for scb in 2 4 16; do echo "scb=$scb"; TSP_SPMM_SCB=$scb python tools/prof_case.py spmm 0.5 5; TSP_SPMM_SCB=$scb python tools/prof_case.py spmm 0.25 5; done
timeout 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:spmm_staged -s 2 -c 1 python tools/prof_case.py spmm 0.5 3 2>&1 | grep -E "dram__|gpu__time|lts__"
timeout 900 python -m pytest tests/test_gpu_spmm.py tests/test_gpu_sddmm.py tests/test_gpu_plans.py tests/test_gpu_golden.py -q 2>&1 | tail -2
