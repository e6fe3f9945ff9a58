#!/bin/bash
# profiles/capture_trace.sh <tag> — plain run first, then ncu --set full (+source) of the two traversal kernels on both bench workloads (one GPU).
set -u
TAG=${1:-r01c}
for W in clumsydragon soup1m; do
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernels --workload $W"
  $CMD > gpurun_out/plain_${W}_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_${W}_$TAG.log; exit 1; }
  tail -1 gpurun_out/plain_${W}_$TAG.log | cut -c1-200
  for K in trace_primary_kernel shadow_shade_kernel; do
    ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -f -o gpurun_out/full_${W}_${K}_$TAG $CMD > gpurun_out/ncu_full_${W}_${K}_$TAG.log 2>&1
    echo "$W $K rc=$?"
  done
done
ls -la gpurun_out/
