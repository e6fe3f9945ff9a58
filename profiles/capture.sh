#!/bin/bash
# profiles/capture.sh <round-tag> — run under gpurun (one GPU).  Per workload: plain run first (must exit 0), then the ncu launch
# list of the same command and `--set full` captures of the kernels.  Outputs land in gpurun_out/; profiles/summarize.py turns
# them into profiles/<tag>_kernels.md + <tag>_launches.csv here (no GPU needed).
set -u
TAG=${1:-r01}
for W in clumsydragon soup1m; do
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload $W"
  $CMD > gpurun_out/plain_${W}_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_${W}_$TAG.log; exit 1; }
  tail -1 gpurun_out/plain_${W}_$TAG.log | cut -c1-300
  ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_${W}_$TAG.csv $CMD > gpurun_out/ncu_list_${W}_$TAG.log 2>&1
  echo "$W launch list rc=$?"
  KS="trace_primary_kernel shadow_shade_kernel"
  if [ $W = clumsydragon ]; then KS="$KS refit_atomic_kernel chain_suffix_kernel morton_leaves_kernel karras_emit_kernel tt_emit_kernel tt_refit_kernel widen_kernel"; fi
  if [ $W = soup1m ]; then KS="$KS refit_atomic_kernel morton_leaves_kernel karras_emit_kernel widen_kernel untile_kernel"; fi
  for K in $KS; do
    ncu --set full --clock-control none --import-source on -k regex:$K -s 6 -c 1 -f -o gpurun_out/full_${W}_${K}_$TAG $CMD > gpurun_out/ncu_full_${W}_${K}_$TAG.log 2>&1
    echo "$W $K rc=$?"
  done
done
# building blocks at 2^24 (sort pass, histogram, scan)
CMD="python profiles/sort_probe.py 24"
$CMD > gpurun_out/plain_sort_$TAG.log 2>&1 || { echo "sort probe failed"; exit 1; }
cat gpurun_out/plain_sort_$TAG.log
for K in onesweep_pass_kernel radix_histogram_kernel scan_lookback_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 8 -c 1 -f -o gpurun_out/full_blocks_${K}_$TAG $CMD > gpurun_out/ncu_full_blocks_${K}_$TAG.log 2>&1
  echo "blocks $K rc=$?"
done
ls gpurun_out | wc -l
