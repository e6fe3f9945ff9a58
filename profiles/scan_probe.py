"""Scan microbench + check (BASELINE.json configs[4]): input splitmix-like small integers, n = 2^e (and ragged sizes); the whole result is
compared with torch.cumsum.  Device time by CUDA events on the context's stream, median of 9, in place and out of place.
usage: scan_probe.py [exponents ...]   (CLRT_LIB_PATH selects a variant build, profiles/variants.sh)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openclraytracer_b200 import api
S = torch.cuda.Stream()
ctx = api.Context(0, stream=S.cuda_stream)
torch.cuda.set_stream(S)
scan = api.PrefixSum(ctx)
exps = [int(a) for a in sys.argv[1:]] or [20, 22, 24, 26, 28]
sizes = [(f"2^{e}", 1 << e) for e in exps] + [("2^24+12345", (1 << 24) + 12345), ("1000003", 1000003), ("777", 777)]
for name, n in sizes:
    a = (torch.arange(n, dtype=torch.int64, device="cuda") * 2654435761 >> 7 & 7).to(torch.int32)
    out = torch.empty_like(a)
    ts = []
    for r in range(10):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda._sleep(400_000)                      # keeps the GPU busy while the host enqueues: launch overhead stays out of the interval
        e0.record(S); scan.computePrefixSum(a.data_ptr(), out.data_ptr(), n); e1.record(S); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ms = float(np.median(ts[1:]))
    want = torch.cumsum(a, 0, dtype=torch.int64).to(torch.int32)
    ok = bool(torch.equal(out, want))
    b = a.clone(); scan.computePrefixSum(b.data_ptr(), b.data_ptr(), n); torch.cuda.synchronize()       # in place
    ok_inplace = bool(torch.equal(b, want))
    odd = torch.zeros(n + 1, dtype=torch.int32, device="cuda"); odd[1:] = a                            # 4-byte aligned only
    scan.computePrefixSum(odd[1:].data_ptr(), odd[1:].data_ptr(), n); torch.cuda.synchronize()
    ok_odd = bool(torch.equal(odd[1:], want))
    print(f"scan {name}: {ms:.4f} ms  {8 * n / ms / 1e6:.0f} GB/s  ok={ok} inplace={ok_inplace} unaligned={ok_odd}", flush=True)
    del a, out, want, b, odd
