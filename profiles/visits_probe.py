import sys; sys.path.insert(0,'.')
import numpy as np
from openclraytracer_b200 import api, scenes, frame
ctx=api.Context(0)
for name in ["soup1m","clumsydragon"]:
    buf = scenes.soup_scene(1_000_000) if name=="soup1m" else scenes.load_bundled(name)
    sc=api.Scene(buf); sc.loadToGPU(ctx); b=api.BVHManager(ctx,sc); b.initialize(); b.initializeFrame(); b.construct()
    cam=frame.orbit_camera(buf,1920,1080,5)
    b.set_stats(True); b.generateContacts(cam); c=b.download_primary_contacts(); b.set_stats(False)
    v=c["pad"][:,1].astype(np.float64).reshape(1080,1920)
    # 8x4 packets
    p=v.reshape(270,4,240,8).transpose(0,2,1,3).reshape(-1,32)
    mx=p.max(axis=1); print(name,"mean visits",v.mean(),"packet eff (sum/32*max)",p.sum()/(32*mx.sum()), "p50/p90/p99/max",np.percentile(v,[50,90,99,100]))
    # ideal with sorted refill (perfect)
    b.free(); sc.free()
