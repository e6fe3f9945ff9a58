#!/bin/bash
# profiles/variants.sh — compile-time variants of the traversal kernels (csrc/trace.cu), built HERE (nvcc, no GPU) and measured under gpurun:
#   bash profiles/variants.sh build name1:"-DCLRT_X=1 -DCLRT_Y=2" name2:...   # -> gpurun_variants/libclrt_<name>.so (git-ignored, travels)
#   gpurun -- 'bash profiles/variants.sh run [workload ...]'                  # profiles/quick_trace.py per variant through CLRT_LIB_PATH
#   VARIANT_OBJS="scan sort" bash profiles/variants.sh build ...                 # variants of other translation units
#   gpurun -- 'bash profiles/variants.sh runcmd <command>'                     # any command per variant
#   gpurun -- 'bash profiles/variants.sh test name'                           # the traversal parity tests against one variant
set -u
cd "$(dirname "$0")/.."
case "${1:-build}" in
  build)
    shift; mkdir -p gpurun_variants
    for spec in "$@"; do
      v="${spec%%:*}"; flags="${spec#*:}"; [ "$flags" = "$spec" ] && flags=""
      ( rm -rf /tmp/clrt_variant_$v; mkdir -p /tmp/clrt_variant_$v
        # only trace.cu depends on the switches: reuse the shipped objects for the rest
        cp openclraytracer_b200/csrc/build/*.o /tmp/clrt_variant_$v/ 2>/dev/null; for o in ${VARIANT_OBJS:-trace}; do rm -f /tmp/clrt_variant_$v/$o.o; done
        make -s -C openclraytracer_b200/csrc EXTRA="$flags" BUILD=/tmp/clrt_variant_$v OUT=$PWD/gpurun_variants/libclrt_$v.so && echo "built $v ($flags)" ) &
    done; wait ;;
  run)
    shift; mkdir -p gpurun_out
    for so in gpurun_variants/libclrt_*.so; do
      CLRT_LIB_PATH=$PWD/$so timeout 300 python profiles/quick_trace.py "$@" 2>&1 | grep -v "^$" | tail -4
    done ;;
  runcmd)                                                                       # any command per variant, e.g. the sort / scan probe
    shift
    for so in gpurun_variants/libclrt_${VARIANT_FILTER:-}*.so; do
      echo "== $so"; CLRT_LIB_PATH=$PWD/$so timeout ${VARIANT_TIMEOUT:-300} bash -c "$*" 2>&1 | tail -8
    done ;;
  test)
    CLRT_LIB_PATH=$PWD/gpurun_variants/libclrt_$2.so timeout 900 python -m pytest tests/test_gpu_bvh.py tests/test_gpu_bounce.py -q -x -m gpu -k "not spheres_10m" 2>&1 | tail -3 ;;
esac
