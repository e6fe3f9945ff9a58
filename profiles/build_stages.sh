#!/bin/bash
# profiles/build_stages.sh [workloads...] — build stage times (bench.py's own measurement) per workload; CLRT_LIB_PATH selects a variant
for w in ${@:-soup1m clumsydragon spheres10m}; do
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --workload $w 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    l=l.strip()
    if l.startswith('{'):
        d=json.loads(l); print('$w', round(d['build_ms'],4), round(d['build_ms_per_Mtri'],4), {k: round(v,4) for k,v in d['build_stage_ms'].items()}, 'value', round(d['value'],1))
"
done
