"""Kernel-only timing of the traversal kernels on both bench scenes (tuning helper; CLRT_LIB_PATH selects the build)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import frame, scenes
for wl in ("clumsydragon", "soup1m"):
    buf, _ = bench.make_scene(wl)
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", 0, 1, 0)
    r.build()
    kt = bench.time_trace_kernels(r, frame.orbit_camera(buf, 1920, 1080, 5), repeats=7)
    print(os.environ.get("CLRT_LIB_PATH", "default").split("/")[-1], wl, "primary %.3f ms  shadow %.3f ms" % (kt["primary"]["ms"], kt["shadow"]["ms"]), flush=True)
    r.close()
