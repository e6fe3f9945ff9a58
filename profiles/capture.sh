#!/bin/bash
# profiles/capture.sh <round-tag> — run under gpurun (one GPU).  Plain run first (must exit 0), then the ncu launch list of the
# same command and `--set full` captures of the top kernels.  Outputs land in gpurun_out/; summaries are copied to profiles/ by hand.
set -u
TAG=${1:-r01}
CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_$TAG.log; exit 1; }
tail -1 gpurun_out/plain_$TAG.log | cut -c1-400
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
for K in trace_primary_kernel trace_rays_kernel onesweep_pass_kernel scan_lookback_kernel refit_atomic_kernel morton_leaves_kernel karras_emit_kernel radix_histogram_kernel shade_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$K -s 4 -c 2 -f -o gpurun_out/full_${K}_$TAG $CMD > gpurun_out/ncu_full_${K}_$TAG.log 2>&1
  echo "$K rc=$?"
done
ls -la gpurun_out/
