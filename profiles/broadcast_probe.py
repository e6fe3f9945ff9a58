"""Multi-GPU probe (torchrun, N >= 2): "the BVH is built once and NCCL-broadcast over NVLink (or rebuilt redundantly, whichever is faster)"
(north star).  Times ncclBroadcast of the bytes the traversal needs (wide nodes + tree-ordered triangles) and of the whole build product
(reference nodes, pairs, TraceNodes, wide nodes, sorted triangles, refpos) against one redundant construct() on every rank.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 profiles/broadcast_probe.py"""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from openclraytracer_b200 import api, frame

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
out = {}
for wl in ("soup1m", "spheres10m"):
    buf, _ = bench.make_scene(wl)
    n = int(np.frombuffer(buf[:96].tobytes(), dtype=np.uint64)[7])
    r = frame.FrameRenderer(buf, 1920, 1080, "BVH", rank, world, local)
    r.accel.enable_stage_timing(True)
    ms = []
    for _ in range(6):
        r.build(); ms.append(r.accel.stage_ms(api.STAGE_TOTAL))
    rebuild = float(np.median(ms[2:]))
    sizes = {"traversal set (64 B wide nodes + 48 B triangles)": 112 * n, "whole build product (396 B per triangle)": 396 * n}
    res = {"triangles": n, "redundant_rebuild_ms": rebuild}
    for name, nbytes in sizes.items():
        t = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        for _ in range(3):
            dist.broadcast(t, 0)
        torch.cuda.synchronize(); dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            dist.broadcast(t, 0)
        e1.record(); torch.cuda.synchronize()
        v = torch.tensor([e0.elapsed_time(e1) / 10], device="cuda"); dist.all_reduce(v, op=dist.ReduceOp.MAX)
        res["ncclBroadcast_ms: " + name] = float(v[0]); res["bytes: " + name] = nbytes
        del t
    out[wl] = res
    if rank == 0:
        print(wl, res, flush=True)
    r.close()
if rank == 0:
    json.dump({"n_gpus": world, "results": out}, open("gpurun_out/broadcast_probe_%d.json" % world, "w"), indent=1)
dist.destroy_process_group()
