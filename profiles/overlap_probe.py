"""Single-GPU probe of the frame-slot pipeline on STRIPES: stripe 0 of nranks rendered back to back through D frame slots on D
streams (no collective) — the per-GPU device throughput an N-GPU run can reach.  CLRT_RAYS_PER_LANE sizes the persistent grids."""
import os, sys
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", os.environ.get("PROBE_CONNECTIONS", "32"))   # before CUDA starts: one hardware queue per stream
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame
D = int(os.environ.get("CLRT_FRAME_DEPTH", "8"))
RANK = int(os.environ.get("PROBE_RANK", "0"))
K = int(os.environ.get("PROBE_BATCH", "1"))          # frames per pair of launches (clrt_render_stripes_batch)
for wl in sys.argv[1:] or ("soup1m",):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    cams = [frame.orbit_camera(buf, 1920, 1080, s) for s in range(32)]
    streams = [torch.cuda.Stream() for _ in range(D)]
    bufs = [torch.zeros(r.cap * K, dtype=torch.int32, device="cuda") for _ in range(D)]
    base = None
    for nr in (1, 2, 4, 8):
        def run(frames):
            if K > 1 and nr > 1:
                for k in range(frames // K):
                    r.accel.render_stripes_batch(streams[k % D].cuda_stream, k % D, [cams[(k * K + j) % 32] for j in range(K)], RANK % nr, nr, bufs[k % D].data_ptr())
                return
            for k in range(frames):
                r.accel.render_stripe_on(streams[k % D].cuda_stream, k % D, cams[k % 32], True, RANK % nr, nr, bufs[k % D].data_ptr())
        run(8); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(r.stream)
        for st in streams: st.wait_event(e0)
        run(40)
        for st in streams: r.stream.wait_stream(st)
        e1.record(r.stream); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 40
        if nr == 1: base = ms
        print(wl, "rank", RANK, "K", K, "D", D, "RPL", os.environ.get("CLRT_RAYS_PER_LANE", "16 (default)"), "nranks", nr, "ms per stripe %.3f  ideal %.3f  efficiency %.2f" % (ms, base / nr, base / nr / ms), flush=True)
    r.close()
