#!/bin/bash
# compute-sanitizer passes over the parity tests of the kernels added or changed last (memcheck: out-of-bounds /
# misaligned accesses; racecheck: shared-memory hazards; synccheck: barrier misuse).  Summaries -> gpurun_out/.
set -u
mkdir -p gpurun_out
SEL='tests/test_zz_no_alloc.py tests/test_gpu_softmax_transpose.py tests/test_gpu_plans.py tests/test_gpu_f16.py'
K='transpose or permute or tiled_view or native_sampled or noalloc'
for tool in memcheck racecheck synccheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 5 --error-exitcode 9 \
    python -m pytest $SEL -q -x -m gpu -k "$K" > gpurun_out/sanitizer_$tool.log 2>&1
  echo "$tool rc=$?"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer_$tool.log | tail -3
done
