"""Single-GPU probe: device time of clrt_render_stripe for stripe 0 of nranks = 1, 2, 4, 8 (how the per-GPU share of a frame
shrinks with the number of ranks, without any collective)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame
for wl in sys.argv[1:] or ("clumsydragon", "soup1m"):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    cam = frame.orbit_camera(buf, 1920, 1080, 5)
    for nr in (1, 2, 4, 8):
        out = []
        for rank in range(min(nr, 2)):
            for _ in range(3):
                r.accel.render_stripe(cam, True, rank, nr, r.stripes[0].data_ptr())
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            with torch.cuda.stream(r.stream):
                e0.record(r.stream)
                for _ in range(10):
                    r.accel.render_stripe(cam, True, rank, nr, r.stripes[0].data_ptr())
                e1.record(r.stream)
            r.stream.synchronize()
            out.append(e0.elapsed_time(e1) / 10)
        print(wl, "nranks", nr, "stripe ms", ["%.3f" % x for x in out], "ideal %.3f" % (out[0] if nr == 1 else base / nr), flush=True)
        if nr == 1: base = out[0]
    r.close()
