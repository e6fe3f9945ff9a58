"""Spheres-10M (BASELINE.json configs[3]) on one B200: LBVH build of 10 M triangles, 3840x2160 primary rays, one incoherent diffuse
bounce per hit through the generic-ray entry.  Device times by CUDA events on the context's stream.  One JSON line."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from openclraytracer_b200 import api, scenes
S = torch.cuda.Stream(); torch.cuda.set_stream(S)
ctx = api.Context(0, stream=S.cuda_stream)
buf = scenes.sphere_field_scene()
scene = api.Scene(buf); scene.loadToGPU(ctx)
bvh = api.BVHManager(ctx, scene); bvh.initialize(); bvh.enable_stage_timing(True)
tot, st = [], []
for _ in range(5):
    bvh.initializeFrame(); bvh.construct()
    tot.append(bvh.stage_ms(api.STAGE_TOTAL)); st.append([bvh.stage_ms(s) for s in (0, 1, 2, 3, 5)])
bvh.enable_stage_timing(False); bvh.initializeFrame(); bvh.construct()
W, H = 3840, 2160; n = W * H
cam = api.demo_camera(buf, W, H)
rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
def timed(fn, reps=5):
    fn(); ctx.finish()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(S)
    for _ in range(reps): fn()
    e1.record(S); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms_p = timed(lambda: bvh.generateContacts(cam))
hits = int((bvh.download_primary_contacts()["nd"][:, 3] > 0).sum())
api.generate_bounce_rays(ctx, cam, bvh.getPrimaryContacts(), n, 20161116, rays.getCLMem())
ms_b = timed(lambda: bvh.generateContacts(rays, contacts, n))
med = np.median(np.array(st[1:]), axis=0)
print(json.dumps({"workload": "Spheres-10M (2000 tessellated spheres, 10e6 triangles), 3840x2160", "build_ms": float(np.median(tot[1:])),
                  "build_ms_per_Mtri": float(np.median(tot[1:])) / 10.0,
                  "stages_ms": dict(zip(("morton", "sort", "emit", "box_pyramid", "trace_structure"), [float(x) for x in med])),
                  "trace_tree": bvh.has_trace_tree(), "primary_ms": ms_p, "primary_Mrays_per_s": n / ms_p / 1e3, "primary_hits": hits,
                  "bounce_ms": ms_b, "bounce_live_rays": hits, "bounce_Mrays_per_s": hits / ms_b / 1e3}))
