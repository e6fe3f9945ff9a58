#!/bin/bash
# profiles/run_multi.sh <N> <tag> [what...] — under `gpurun --gpus N`: multi-GPU measurements on N GPUs of one box.
#   what: exchange (parity of all exchanges on every rank + their frame rates), bench (soup1m, two kernels per frame),
#         bench1k (the one-kernel frame), spheres (configs[3] at N GPUs), broadcast (ncclBroadcast vs redundant rebuild)
N=$1; TAG=$2; shift 2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port"
line() { python - "$1" <<'PY'
import json, sys
try:
    d = json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print("%s: %.1f %s  %.4f ms/step  e2e %.1f (%.4f ms)  gpus %d" % (sys.argv[1], d["value"], d["unit"], d["ms_per_step"], d["e2e"]["value"], d["e2e"]["ms_per_step"], d["n_gpus"]))
except Exception as e:
    print(sys.argv[1], "no line:", e)
PY
}
for what in "$@"; do
  case $what in
    exchange) PROBE_STEPS=200 PROBE_MODES=${PROBE_MODES:-nccl,fused,fused-split} timeout 600 $TR 29511 profiles/exchange_probe.py soup1m 2>&1 | grep -v "^\[W\|^W0\|^$" | tail -8
              cp gpurun_out/exchange_probe_$N.json gpurun_out/exchange_${N}gpu_$TAG.json 2>/dev/null ;;
    bench)    timeout 600 $TR 29513 bench.py --gpus $N --steps 400 --warmup 20 --no-kernels > gpurun_out/bench_soup1m_${N}gpu_$TAG.json 2> gpurun_out/bench_soup1m_${N}gpu_$TAG.err; line gpurun_out/bench_soup1m_${N}gpu_$TAG.json ;;
    bench1k)  CLRT_FRAME_KERNELS=1 timeout 600 $TR 29514 bench.py --gpus $N --steps 400 --warmup 20 --no-kernels > gpurun_out/bench_soup1m_1k_${N}gpu_$TAG.json 2> gpurun_out/bench_soup1m_1k_${N}gpu_$TAG.err; line gpurun_out/bench_soup1m_1k_${N}gpu_$TAG.json ;;
    benchnosplit) CLRT_SPLIT_EXCHANGE=0 timeout 600 $TR 29515 bench.py --gpus $N --steps 400 --warmup 20 --no-kernels > gpurun_out/bench_soup1m_nosplit_${N}gpu_$TAG.json 2> gpurun_out/bench_soup1m_nosplit_${N}gpu_$TAG.err; line gpurun_out/bench_soup1m_nosplit_${N}gpu_$TAG.json ;;
    cfg:*)    # cfg:<label>:<ENV=val,ENV=val,...> — bench.py soup1m under the given environment
              label=$(echo $what | cut -d: -f2); envs=$(echo $what | cut -d: -f3 | tr "," " ")
              env $envs timeout 600 $TR 29518 bench.py --gpus $N --steps 300 --warmup 20 --no-kernels > gpurun_out/bench_soup1m_${label}_${N}gpu_$TAG.json 2> gpurun_out/bench_soup1m_${label}_${N}gpu_$TAG.err; line gpurun_out/bench_soup1m_${label}_${N}gpu_$TAG.json ;;
    bench2k16) CLRT_FRAME_KERNELS=2 CLRT_FRAME_DEPTH=16 timeout 600 $TR 29517 bench.py --gpus $N --steps 400 --warmup 20 --no-kernels > gpurun_out/bench_soup1m_2k16_${N}gpu_$TAG.json 2> gpurun_out/bench_soup1m_2k16_${N}gpu_$TAG.err; line gpurun_out/bench_soup1m_2k16_${N}gpu_$TAG.json ;;
    spheres)  timeout 900 $TR 29516 bench.py --gpus $N --workload spheres10m --steps 40 --warmup 5 --no-kernels > gpurun_out/bench_spheres10m_${N}gpu_$TAG.json 2> gpurun_out/bench_spheres10m_${N}gpu_$TAG.err; line gpurun_out/bench_spheres10m_${N}gpu_$TAG.json ;;
    broadcast) timeout 600 $TR 29512 profiles/broadcast_probe.py 2>&1 | grep -v "^\[W\|^W0\|^$" | tail -3; cp gpurun_out/broadcast_probe_$N.json gpurun_out/broadcast_${N}gpu_$TAG.json 2>/dev/null ;;
  esac
done
