#!/bin/bash
# profiles/capture_soup.sh <tag> — ncu --set full of the two traversal kernels on the Soup-1M workload (one GPU).
set -u
TAG=${1:-r01b}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-kernels --workload soup1m"
$CMD > gpurun_out/plain_soup_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_soup_$TAG.log; exit 1; }
tail -1 gpurun_out/plain_soup_$TAG.log | cut -c1-300
for K in trace_primary_kernel shadow_shade_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -f -o gpurun_out/full_soup_${K}_$TAG $CMD > gpurun_out/ncu_full_soup_${K}_$TAG.log 2>&1
  echo "$K rc=$?"
done
