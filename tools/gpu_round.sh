#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench, ncu launch lists (each ncu pass only after its command exited 0 without ncu).
set -u
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -3 gpurun_out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -2
timeout 900 python bench.py > gpurun_out/r02_bench_n1.json 2> gpurun_out/bench.err; rc=$?; echo "bench rc=$rc"
tail -c 400 gpurun_out/bench.err
if [ $rc -eq 0 ]; then
  timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r02_launches.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-ref-kernel > gpurun_out/ncu_bench.log 2>&1; echo "ncu rc=$?"
  python tools/launch_shares.py gpurun_out/r02_launches.csv > gpurun_out/r02_launch_shares.txt 2>&1; head -30 gpurun_out/r02_launch_shares.txt
fi
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"
timeout 600 python tools/train_dp_transformer.py --graph 0 --steps 5 > gpurun_out/r02_c5_eager_n1.json 2> gpurun_out/c5.err; echo "c5 eager rc=$?"; cat gpurun_out/r02_c5_eager_n1.json
timeout 600 python tools/train_dp_transformer.py --graph 1 --steps 10 > gpurun_out/r02_c5_graph_n1.json 2>> gpurun_out/c5.err; echo "c5 graph rc=$?"; cat gpurun_out/r02_c5_graph_n1.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r02_c5_launches.csv \
  python tools/train_dp_transformer.py --graph 0 --steps 2 --warmup 2 > gpurun_out/ncu_c5.log 2>&1; echo "ncu c5 rc=$?"
python tools/launch_shares.py gpurun_out/r02_c5_launches.csv 0.8 1.0 > gpurun_out/r02_c5_launch_shares_eager_step.txt 2>&1; head -40 gpurun_out/r02_c5_launch_shares_eager_step.txt
timeout 600 python tools/bench_configs.py --config all > gpurun_out/r02_configs_n1.jsonl 2> gpurun_out/configs.err; echo "configs rc=$?"
