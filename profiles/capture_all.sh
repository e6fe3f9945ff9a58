#!/bin/bash
# profiles/capture_all.sh <tag> — under gpurun (one GPU): `ncu --set full` over profiles/all_kernels_probe.py, i.e. a capture of EVERY
# kernel of the library at bench sizes (plain run first, must exit 0).  The report is turned into CSV on the box (the raw page: one row
# per launch, every metric) so that only a few MB travel back; profiles/summarize_all.py <tag> then writes profiles/<tag>_all_kernels.md
# and profiles/<tag>_traffic.json here (no GPU needed).
# With a second argument (a kernel-name regex) only those kernels are profiled and the CSV is gpurun_out/all_<tag>_part.csv, which
# summarize_all.py merges with the full pass — for kernels a timed-out full pass did not reach.
set -u
TAG=${1:-r02}
CMD="python profiles/all_kernels_probe.py"
if [ -n "${2:-}" ]; then
  timeout 600 ncu --set full --clock-control none -k "regex:$2" -f -o /tmp/all_${TAG}_part $CMD > gpurun_out/ncu_all_${TAG}_part.log 2>&1
  echo "ncu (filtered) rc=$?"
  ncu -i /tmp/all_${TAG}_part.ncu-rep --page raw --csv > gpurun_out/all_${TAG}_part.csv 2>/dev/null
  ls -la gpurun_out/all_${TAG}_part.csv | awk '{print $5, $9}'
  exit 0
fi
timeout 400 $CMD > gpurun_out/plain_all_$TAG.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_all_$TAG.log; exit 1; }
tail -1 gpurun_out/plain_all_$TAG.log
timeout 1500 ncu --set full --clock-control none -f -o /tmp/all_$TAG $CMD > gpurun_out/ncu_all_$TAG.log 2>&1
echo "ncu rc=$?"
ncu -i /tmp/all_$TAG.ncu-rep --page raw --csv > gpurun_out/all_$TAG.csv 2>/dev/null
ls -la gpurun_out/all_$TAG.csv /tmp/all_$TAG.ncu-rep | awk '{print $5, $9}'
