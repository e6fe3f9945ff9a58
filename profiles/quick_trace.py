"""Kernel-only timing of the traversal kernels on the bench scenes + the pipelined frame time (tuning helper; CLRT_LIB_PATH selects the
build).  usage: quick_trace.py [workload ...]"""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame, scenes
tag = os.environ.get("CLRT_LIB_PATH", "default").split("/")[-1].replace("libclrt_", "").replace(".so", "")
for wl in (sys.argv[1:] or ["clumsydragon", "soup1m"]):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    kt = bench.time_trace_kernels(r, frame.orbit_camera(buf, 1920, 1080, 5), repeats=7)
    cams = [frame.orbit_camera(buf, 1920, 1080, s) for s in range(40)]
    for c in cams[:8]:
        r.render_device(c, True)
    r.finish()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(r.stream)
    for c in cams[8:]:
        r.render_device(c, True)
    r.join(); e1.record(r.stream); e1.synchronize()
    r.accel.check()
    print("[%s] %s primary %.3f ms (%.1f visits %.2f tests)  shadow %.3f ms (%.1f visits %.2f tests)  frame %.3f ms" % (
        tag, wl, kt["primary"]["ms"], kt["primary"]["inner_per_ray"], kt["primary"]["tris_per_ray"],
        kt["shadow"]["ms"], kt["shadow"]["inner_per_ray"], kt["shadow"]["tris_per_ray"], e0.elapsed_time(e1) / 32), flush=True)
    r.close()
