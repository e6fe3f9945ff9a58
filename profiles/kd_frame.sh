#!/bin/bash
# profiles/grid_frame.sh — one line per run: the Two-Level Grid frame (bench.py --accel KD, corrected walk) on Soup-1M; CLRT_LIB_PATH selects a variant
timeout 200 python bench.py --accel KD --steps 20 --warmup 3 --no-kernels --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    l = l.strip()
    if l.startswith('{'):
        d = json.loads(l); print('kd frame %.3f ms  %.1f Mrays/s  e2e %.1f' % (d['ms_per_step'], d['value'], d['e2e']['value']), {k: round(v['ms'], 3) for k, v in d.get('trace_kernels', {}).items()})
"
