"""Multi-GPU probe (run under torchrun, N >= 2): the three frame exchanges of frame.FrameRenderer — NCCL all-gather + untile,
"push" and "fused" over peer memory — must give every rank the same raster, bit for bit; then their frame rates.
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 profiles/exchange_probe.py soup1m
"""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import bench
from openclraytracer_b200 import frame

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
RES = (1920, 1080)
CHECK, STEPS = 10, int(os.environ.get("PROBE_STEPS", "200"))
out = {}
for wl in sys.argv[1:] or ("soup1m",):
    buf, _ = bench.make_scene(wl)
    cams = [frame.orbit_camera(buf, RES[0], RES[1], i) for i in range(64)]
    golden = None
    for mode in os.environ.get("PROBE_MODES", "nccl,push,fused,fused-split").split(","):
        os.environ["CLRT_EXCHANGE"] = mode.split("-")[0]
        os.environ["CLRT_SPLIT_EXCHANGE"] = "1" if mode.endswith("-split") else "0"      # waiting half of the exchange on its own stream
        r = frame.FrameRenderer(buf, RES[0], RES[1], "BVH", rank, world, local)
        assert r.exchange == mode.split("-")[0]
        r.build()
        # parity: frame by frame on every rank (the frames of a pipeline of DEPTH slots, twice around)
        frames = []
        for i in range(CHECK):
            r.render_device(cams[i], True)
            r.finish()
            frames.append(r.raster.cpu().numpy().copy())
        host = [r.render_host(cams[i], True).numpy().copy() for i in range(2)]
        if rank == 0:
            assert np.array_equal(host[0], frames[0]) and np.array_equal(host[1], frames[1]), mode + ": host read-back differs"
        if golden is None:
            golden = frames
            assert len({f.tobytes() for f in frames}) == CHECK, "frames of a moving camera must differ"
        else:
            for i in range(CHECK):
                bad = int((frames[i] != golden[i]).sum())
                assert bad == 0, "%s: frame %d differs from the NCCL exchange in %d pixels on rank %d" % (mode, i, bad, rank)
        # pipelined: no host wait between frames; the last DEPTH frames are still checked afterwards
        for _ in range(2):
            for i in range(16): r.render_device(cams[i % 64], True)
            r.finish()
        dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(r.stream)
        t0 = time.perf_counter()
        for i in range(STEPS): r.render_device(cams[i % 64], True)
        r.join(); e1.record(r.stream)
        host_ms = (time.perf_counter() - t0) * 1e3 / STEPS
        r.finish()
        t = torch.tensor([e0.elapsed_time(e1) / STEPS], device="cuda"); dist.all_reduce(t, op=dist.ReduceOp.MAX)
        last = r.raster.cpu().numpy().copy()
        r.render_device(cams[(STEPS - 1) % 64], True); r.finish()
        assert np.array_equal(last, r.raster.cpu().numpy()), mode + ": pipelined frame differs from the same frame rendered alone"
        dist.barrier(); torch.cuda.synchronize()
        e0.record(r.stream)
        for i in range(STEPS): r.render_host_async(cams[i % 64], True)
        r.join(); e1.record(r.stream); r.finish()
        t2 = torch.tensor([e0.elapsed_time(e1) / STEPS], device="cuda"); dist.all_reduce(t2, op=dist.ReduceOp.MAX)
        if rank == 0:
            out.setdefault(wl, {})[mode] = {"device_ms_per_frame": float(t[0]), "e2e_ms_per_frame": float(t2[0]), "host_submit_ms_per_frame": host_ms}
            print(wl, mode, out[wl][mode], flush=True)
        r.close()
        dist.barrier()
if rank == 0:
    os.makedirs("gpurun_out", exist_ok=True)
    json.dump({"n_gpus": world, "resolution": list(RES), "steps": STEPS, "parity": "all modes bit-identical on every rank", "results": out},
              open("gpurun_out/exchange_probe_%d.json" % world, "w"), indent=1)
dist.destroy_process_group()
