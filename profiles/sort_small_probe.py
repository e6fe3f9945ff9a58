"""Sort at the build's sizes (30-bit keys through the 32-bit entry point): 2^17 .. 2^21 and 1 000 000 pairs, median of 15, GPU kept busy ahead
of the first event so that launch overhead stays out.  CLRT_LIB_PATH selects a variant (profiles/variants.sh)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openclraytracer_b200 import api
S = torch.cuda.Stream(); torch.cuda.set_stream(S)
ctx = api.Context(0, stream=S.cuda_stream)
sorter = api.BitonicSort(ctx)
for n in (1 << 17, 1 << 18, 1 << 19, 1_000_000, 1 << 20, 1 << 21, 1 << 22):
    g = torch.Generator(device="cuda"); g.manual_seed(n)
    keys = torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device="cuda", generator=g)
    pairs0 = torch.stack([keys, torch.arange(n, dtype=torch.int32, device="cuda")], dim=1).contiguous()
    pairs = pairs0.clone(); ts = []
    for r in range(16):
        pairs.copy_(pairs0)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(300_000)
        e0.record(S); sorter.sort(pairs.data_ptr(), n); e1.record(S); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    k = pairs[:, 0]
    print(f"sort {n}: {np.median(ts[1:]) * 1e3:.1f} us  sorted={bool((k[1:] >= k[:-1]).all())}", flush=True)
