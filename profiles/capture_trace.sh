#!/bin/bash
# profiles/capture_trace.sh <tag> [workload] — under gpurun (one GPU): `ncu --set full` + source of ONE launch of each traversal kernel
# of the frame (profiles/quick_trace.py <workload>; plain run first).  CLRT_LIB_PATH selects a variant build.
set -u
TAG=${1:-r02}; W=${2:-soup1m}
CMD="python profiles/quick_trace.py $W"
$CMD > gpurun_out/plain_trace_${W}_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_trace_${W}_$TAG.log; exit 1; }
tail -1 gpurun_out/plain_trace_${W}_$TAG.log
for K in trace_primary_kernel shadow_shade_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 1 -f -o gpurun_out/full_${W}_${K}_$TAG $CMD > gpurun_out/ncu_full_${W}_${K}_$TAG.log 2>&1
  echo "$W $K rc=$?"
done
