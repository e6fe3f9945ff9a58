import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from openclraytracer_b200 import api, scenes
ctx = api.Context(0)
buf = scenes.soup_scene(1_000_000)
scene = api.Scene(buf); scene.loadToGPU(ctx)
bvh = api.BVHManager(ctx, scene); bvh.initialize(); bvh.enable_stage_timing(True)
st = []
for i in range(8):
    bvh.initializeFrame(); bvh.construct()
    st.append([bvh.stage_ms(s) for s in (0, 1, 2, 3, 4, 5)])
print("morton sort emit pyramid total wide:", np.round(np.median(np.array(st[2:]), axis=0), 4))
