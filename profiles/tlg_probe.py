"""Two-Level Grid on the bench scenes (BASELINE.json configs[2]: "LBVH vs Two-Level Grid build + traversal"): build stage times
and one 1080p primary + shadow frame through FrameRenderer(accel="GRID"), next to the LBVH numbers.  One JSON line per scene."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import bench
from openclraytracer_b200 import api, frame
for wl in sys.argv[1:] or ("clumsydragon", "soup1m"):
    buf, desc = bench.make_scene(wl)
    out = {"workload": desc}
    for acc in ("BVH", "GRID"):
        r = frame.FrameRenderer(buf, 1920, 1080, acc, 0, 1, 0)
        r.accel.enable_stage_timing(True)
        tot, st = [], []
        for _ in range(5):
            r.build()
            tot.append(r.accel.stage_ms(api.STAGE_TOTAL))
            st.append([r.accel.stage_ms(s) for s in ((0, 1, 2, 3, 5) if acc == "BVH" else (8, 9, 10, 11, 12))])
        r.accel.enable_stage_timing(False); r.build()
        cams = [frame.orbit_camera(buf, 1920, 1080, s) for s in range(13)]
        rays = sum(r.render_device(c, True, want_rays=True) for c in cams[3:])
        r.finish()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(r.stream)
        for c in cams[3:]:
            r.render_device(c, True)
        r.join(); e1.record(r.stream); r.finish()
        ms = e0.elapsed_time(e1) / 10
        names = ("morton", "sort", "emit", "box_pyramid", "trace_structure") if acc == "BVH" else ("count_top", "top_pairs", "leaf_cells", "leaf_count", "leaf_pairs")
        out[acc] = {"build_ms": float(np.median(tot[1:])), "stages_ms": dict(zip(names, [float(x) for x in np.median(np.array(st[1:]), axis=0)])),
                    "frame_ms": ms, "Mrays_per_s": rays / 10 / ms / 1e3}
        if acc == "GRID":
            out[acc]["sizes(top_cells, top_pairs, leaf_cells, leaf_pairs)"] = [int(x) for x in r.accel.sizes()]
        r.close()
    print(json.dumps(out), flush=True)
