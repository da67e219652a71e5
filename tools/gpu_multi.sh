#!/bin/bash
# Multi-GPU pass on one box: bench.py (headline + scale_cases) at N ranks, then the C5 training step under a CUDA graph.
set -u
mkdir -p gpurun_out
N=${1:-8}
nvidia-smi topo -m > gpurun_out/r02_topo_n$N.txt 2>&1
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench rc=$?"
tail -c 600 gpurun_out/bench_n$N.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/train_dp_transformer.py --graph 1 --steps 10 > gpurun_out/r02_c5_graph_n$N.json 2> gpurun_out/c5_n$N.err; echo "c5 rc=$?"
cat gpurun_out/r02_c5_graph_n$N.json
python - $N <<'PY'
import json,sys
d=json.load(open(f'gpurun_out/r02_bench_n{sys.argv[1]}.json'))
print(d['n_gpus'], d['value'], d['ms_per_step'], d['gpu_launches'], d['e2e']['value'], d['e2e']['ms_per_step'], d['e2e']['host_copy_ceiling'])
print(json.dumps({k:{kk:vv for kk,vv in v.items() if kk!='what'} for k,v in d['scale_cases'].items()}, indent=1))
PY
