"""Sort / scan building-block microbench (BASELINE.json configs[4]): (key = splitmix64(i) & 0x3FFFFFFF, value = i) pairs,
n = 2^20 .. 2^28; scan input splitmix64(i) & 7.  Device time by CUDA events on the context's stream, median of 5."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openclraytracer_b200 import api, scenes
S = torch.cuda.Stream()
ctx = api.Context(0, stream=S.cuda_stream)
torch.cuda.set_stream(S)
sorter = api.BitonicSort(ctx); scan = api.PrefixSum(ctx)
exps = [int(a) for a in sys.argv[1:]] or [20, 22, 24, 26, 28]
for e in exps:
    n = 1 << e
    i = torch.arange(n, dtype=torch.int64, device="cuda")
    z = i + 0x9E3779B97F4A7C15 - (1 << 64) if False else i
    # splitmix64 on the device (int64 wraps like uint64)
    def mix(x):
        x = x + (-7046029254386353131)
        x = (x ^ ((x >> 30) & 0x3FFFFFFFF)) * (-4658895280553007687)
        x = (x ^ ((x >> 27) & 0x1FFFFFFFFF)) * (-7723592293110705685)
        return x ^ ((x >> 31) & 0x1FFFFFFFF)
    keys = (mix(i) & 0x3FFFFFFF).to(torch.int32)
    pairs0 = torch.stack([keys, i.to(torch.int32)], dim=1).contiguous()
    del i
    pairs = pairs0.clone()
    ts = []
    for r in range(6):
        pairs.copy_(pairs0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(S); sorter.sort(pairs.data_ptr(), n); e1.record(S); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts[1:]))
    k = pairs[:, 0].to(torch.int64) & 0xFFFFFFFF
    ok = bool((k[1:] >= k[:-1]).all())
    print(f"sort 2^{e}: {ms:.3f} ms  {n / ms / 1e6:.2f} Gpairs/s  {16 * n / ms / 1e6:.0f} GB/s algorithmic  sorted={ok}", flush=True)
    del pairs, pairs0, keys, k
    a = (torch.arange(n, dtype=torch.int64, device="cuda") * 2654435761 >> 7 & 7).to(torch.int32)
    out = torch.empty_like(a)
    ts = []
    for r in range(6):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(S); scan.computePrefixSum(a.data_ptr(), out.data_ptr(), n); e1.record(S); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts[1:]))
    ok = bool(out[-1].item() == int(a.sum().item()) & 0xFFFFFFFF) if e <= 28 else True
    print(f"scan 2^{e}: {ms:.3f} ms  {8 * n / ms / 1e6:.0f} GB/s  last_ok={ok}", flush=True)
    del a, out
