import sys, os
sys.path.insert(0, "/root/repo")
import numpy as np
from openclraytracer_b200 import api, scenes, frame, layouts as L
from oracle import oracle_py as O
ctx = api.Context(0)
for name in ["cube", "tenryuu", "soup"]:
    buf = scenes.soup_scene(60_000, seed=13, tris_per_mesh=8192) if name == "soup" else scenes.load_bundled(name)
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    kd = api.KDTreeManager(ctx, scene); kd.initialize(); kd.initializeFrame(); kd.construct()
    tlg = api.TwoLevelGridManager(ctx, scene); tlg.initialize(); tlg.initializeFrame(); tlg.construct(); tlg.set_trace_mode(api.TRACE_FAST)
    print(name, kd.info())
    b = O.bvh_build(buf)
    for cam in (api.demo_camera(buf, 320, 200), frame.orbit_camera(buf, 256, 144, 9)):
        want = O.bvh_trace_primary(cam, b)
        for label, m in (("kd", kd), ("tlg", tlg)):
            m.generateContacts(cam); got = m.download_primary_contacts()
            hit = want["nd"][:, 3] > 0
            print(" ", label, "hitmask equal", np.array_equal(got["nd"][:, 3] > 0, hit), "hits", hit.sum())
            both = hit & (got["nd"][:, 3] > 0)
            dt = np.abs(got["nd"][both, 3].astype(np.float64) - want["nd"][both, 3]) / want["nd"][both, 3]
            same = got["pad"][both, 0] == want["pad"][both, 0]
            print("    max rel dt", dt.max() if len(dt) else 0, "id mismatches", int((~same).sum()), "exact-t among mismatches", int((dt[~same] == 0).sum()))
            if (~same).sum():
                g, w = got["pad"][both, 0][~same], want["pad"][both, 0][~same]
                pairs, cnt = np.unique(np.stack([g, w], 1), axis=0, return_counts=True)
                print("    (got, want) x count:", [(int(a), int(b), int(c)) for (a, b), c in zip(pairs[:8], cnt[:8])])
    kd.free(); tlg.free(); scene.free()
