#!/bin/bash
# profiles/final_run.sh <tag> — under gpurun (one GPU): the whole single-GPU record of a round with the shipped library —
# GPU tests, every bench line, the ncu launch list of the default bench command, the --set full capture of every kernel.
set -u
TAG=${1:-r02}
timeout 1200 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
bash profiles/run_benches.sh $TAG
CMD="python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-kernels"
timeout 300 $CMD > gpurun_out/plain_list_$TAG.log 2>&1 && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/launches_soup1m_$TAG.csv $CMD > gpurun_out/ncu_list_$TAG.log 2>&1
echo "launch list rc=$?"
bash profiles/capture_all.sh $TAG
