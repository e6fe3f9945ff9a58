#!/bin/bash
# profiles/run_benches.sh <tag> — under gpurun (one GPU): the bench lines of every workload / structure + the reference arm.
TAG=${1:-r02}
run() { name=$1; shift; timeout 420 python bench.py "$@" > gpurun_out/bench_${name}_$TAG.json 2> gpurun_out/bench_${name}_$TAG.err || { echo "$name failed"; tail -3 gpurun_out/bench_${name}_$TAG.err; }
        python - "$name" "$TAG" <<'PY'
import json, sys
n, t = sys.argv[1:3]
try:
    d = json.loads(open("gpurun_out/bench_%s_%s.json" % (n, t)).read().strip().splitlines()[-1])
    print("%-14s %9.1f %s  %.3f ms/step  e2e %.1f  build %.3f ms/Mtri" % (n, d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d.get("build_ms_per_Mtri", 0)))
except Exception as e:
    print(n, "no line:", e)
PY
}
run soup1m --steps 200 --warmup 10
run clumsydragon --workload clumsydragon --steps 200 --warmup 10 --no-kernels
run grid --accel GRID --steps 40 --warmup 5 --no-kernels --no-cpu-baseline
run kd --accel KD --steps 40 --warmup 5 --no-kernels --no-cpu-baseline
run spheres10m --workload spheres10m --steps 30 --warmup 5 --no-kernels
run reference --impl reference --steps 10 --warmup 1
