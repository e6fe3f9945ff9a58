"""Single-GPU probe: where does a stripe's time go when the stripe shrinks?  primary-only vs primary+shadow stripes, and whole
frames at lower resolutions (same rays-per-GPU as a stripe, but spatially dense)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame
def timeit(r, fn, reps=10):
    for _ in range(3): fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(r.stream):
        e0.record(r.stream)
        for _ in range(reps): fn()
        e1.record(r.stream)
    r.stream.synchronize()
    return e0.elapsed_time(e1) / reps
for wl in sys.argv[1:] or ("soup1m",):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    cam = frame.orbit_camera(buf, 1920, 1080, 5)
    for nr in (1, 2, 8):
        p = timeit(r, lambda: r.accel.render_stripe(cam, False, 0, nr, r.stripes[0].data_ptr()))
        ps = timeit(r, lambda: r.accel.render_stripe(cam, True, 0, nr, r.stripes[0].data_ptr()))
        print(wl, "stripe 0 of", nr, "primary+shade %.3f ms   primary+shadow-shade %.3f ms" % (p, ps), flush=True)
    for res in ((1920, 1080), (960, 540), (680, 382), (480, 270)):
        c = frame.orbit_camera(buf, res[0], res[1], 5)
        p = timeit(r, lambda: r.accel.render_stripe(c, False, 0, 1, r.stripes[0].data_ptr()))
        ps = timeit(r, lambda: r.accel.render_stripe(c, True, 0, 1, r.stripes[0].data_ptr()))
        print(wl, "full frame", res, "primary+shade %.3f ms   primary+shadow-shade %.3f ms" % (p, ps), flush=True)
    r.close()
