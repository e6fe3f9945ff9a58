"""How long does the HOST take to submit one frame (no device wait)?  Compared with the device time per frame it tells whether a
workload is bound by submission.  One GPU."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame
for wl in sys.argv[1:] or ("clumsydragon",):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    cams = [frame.orbit_camera(buf, 1920, 1080, s) for s in range(64)]
    for c in cams[:8]: r.render_device(c, True)
    r.finish()
    # tiny frames: the device is never the bottleneck, so wall time per call = host submission cost
    small = frame.FrameRenderer(buf, 64, 32, "BVH", 0, 1, 0); small.build()
    sc = [frame.orbit_camera(buf, 64, 32, s) for s in range(64)]
    for c in sc[:8]: small.render_device(c, True)
    small.finish()
    t0 = time.perf_counter()
    for k in range(400): small.render_device(sc[k % 64], True)
    t1 = time.perf_counter(); small.finish()
    t2 = time.perf_counter()
    for k in range(400): small.render_host_async(sc[k % 64], True)
    t3 = time.perf_counter(); small.finish()
    # full frames: wall per frame including the device
    t4 = time.perf_counter()
    for k in range(400): r.render_device(cams[k % 64], True)
    r.finish(); t5 = time.perf_counter()
    print(wl, "host submit us/frame: device-resident %.1f, e2e %.1f;  1080p wall us/frame %.1f" % ((t1 - t0) / 400 * 1e6, (t3 - t2) / 400 * 1e6, (t5 - t4) / 400 * 1e6), flush=True)
    r.close(); small.close()
