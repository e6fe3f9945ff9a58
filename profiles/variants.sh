#!/bin/bash
# profiles/variants.sh — build the prepared compile-time experiments of csrc/trace.cu HERE (nvcc, no GPU), then run under gpurun:
#   bash profiles/variants.sh build            # -> gpurun_variants/libclrt_<name>.so (git-ignored, travels to the GPU box)
#   gpurun -- 'bash profiles/variants.sh run'  # parity tests + short bench per variant through CLRT_LIB_PATH
# Variants: base (shipped flags), top (CLRT_STACK_TOP_REG=1), ahead (CLRT_PACKET_AHEAD=1), both.
set -u
cd "$(dirname "$0")/.."
declare -A FLAGS=( [base]="" [top]="-DCLRT_STACK_TOP_REG=1" [ahead]="-DCLRT_PACKET_AHEAD=1" [both]="-DCLRT_STACK_TOP_REG=1 -DCLRT_PACKET_AHEAD=1" )
case "${1:-build}" in
  build)
    mkdir -p gpurun_variants
    for v in base top ahead both; do
      make -s -C openclraytracer_b200/csrc -j8 EXTRA="${FLAGS[$v]}" BUILD=/tmp/clrt_variant_$v OUT=$PWD/gpurun_variants/libclrt_$v.so || exit 1
      echo "built gpurun_variants/libclrt_$v.so (${FLAGS[$v]:-shipped flags})"
    done ;;
  run)
    mkdir -p gpurun_out
    for v in base top ahead both; do
      export CLRT_LIB_PATH=$PWD/gpurun_variants/libclrt_$v.so
      timeout 300 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_bounce.py -q -x -m gpu -k "not spheres_10m and not full_size" 2>&1 | tail -1 | sed "s/^/[$v] /"
      for w in soup1m clumsydragon; do
        timeout 200 python bench.py --workload $w --steps 20 --warmup 5 --no-cpu-baseline --no-kernels 2>/dev/null | tail -1 > gpurun_out/variant_${v}_$w.json
        python - "$v" "$w" <<'PY'
import json, sys
v, w = sys.argv[1:3]
try:
    d = json.loads(open("gpurun_out/variant_%s_%s.json" % (v, w)).read())
    print("[%s] %s: %.0f Mrays/s, %.4f ms/frame; primary %.4f ms, shadow %.4f ms" % (v, w, d["value"], d["ms_per_step"], d["trace_kernels"]["primary"]["ms"], d["trace_kernels"]["shadow"]["ms"]))
except Exception as e:
    print("[%s] %s: no line (%s)" % (v, w, e))
PY
      done
    done ;;
esac
