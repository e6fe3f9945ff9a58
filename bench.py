#!/usr/bin/env python
"""bench.py — the headline benchmark of BASELINE.json on N B200s of one box.

metric   : Mrays/s (primary + shadow rays, 1920x1080), whole job.  Also reported: BVH build ms per M triangles.
workload : default `soup1m` = BASELINE.json configs[2], the configuration the metric is quoted on at 1/2/4/8 B200 (synthetic
           1M-triangle random soup, 46 sub-meshes); `--workload clumsydragon` = configs[1], the largest bundled TestModels mesh
           present (50 032 triangles; Skull / swfighter are missing blobs).  LBVH build once, then per step ONE 1080p frame:
           primary rays, one shadow ray per hit per light (scene light + the reference's camera light), Blinn-Phong shading.
step     : one frame of the moving demo camera.  N > 1: the frame is sharded in image space (tile stripes), the
           RGBA8 stripes are all-gathered — total work fixed => "scaling": "strong".
value    : rays traced by all ranks / device time (CUDA events on the launching stream, max over ranks), with the
           camera and the built structure already resident.
e2e      : the same through FrameRenderer.render_host_async (what a user calls; frames in flight): host camera in, pinned-host RGBA8 raster out.
roofline : the dominant kernel of the step (BVH traversal): algorithmic bytes = wide-node + triangle + contact bytes its
           rays touch (counted by the kernel's statistics mode),
           against the measured HBM copy peak — traversal is latency/divergence bound, so this fraction explains
           rather than bounds; the HBM-bound build kernels are reported per kernel under "kernels".
cpu_baseline / --impl reference : the reference's own CPU code for the path (oracle/_ref/libclrt_ref.so = its headers
           compiled in place; else the oracle port) on the host cores, bounded sample.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Mrays/s (primary+shadow, 1080p)"
UNIT = "Mrays/s"
RES = (1920, 1080)
DATA = {"clumsydragon": "bundled mesh, synthetic camera path", "soup1m": "synthetic"}


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def make_scene(workload):
    from openclraytracer_b200 import scenes
    if workload == "clumsydragon":
        return scenes.load_bundled("clumsydragon"), "ClumsyDragon (50032 tris, bundled TestModels) + LIGHT 0 10 -330 1000"
    if workload == "soup1m":
        return scenes.soup_scene(1_000_000), "Soup-1M (1e6 random triangles, 46 sub-meshes, seed 20161116)"
    raise SystemExit(f"unknown workload {workload}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        self.thread.join(timeout=2)
        sm, smax, reasons = [], None, set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); smax = float(r[1])
            except (ValueError, IndexError):
                continue
            for name, v in zip(["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"], r[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons), "samples": len(sm)}


# ---------------------------------------------------------------------------------------------------------------
class CpuReference:
    """The reference's CPU implementation of the step: primary rays through the reference's own bvh_generate_contact,
    one shadow ray per hit per light through its generic-ray path (the reference has no shadow rays; they are
    generated by the oracle, untimed), cpuShadeBlinnPhong.  oracle/_ref/libclrt_ref.so (the reference's headers compiled
    in place, OpenMP over work-items) when present ("reference"), else the oracle port ("port")."""

    def __init__(self, scene_buf):
        from oracle import oracle_py as O
        O.build(("oracle",))
        self.O, self.scene = O, scene_buf
        self.kind = "reference" if O.have_ref() else "port"
        # all the host cores this process may run on — torchrun exports OMP_NUM_THREADS=1 to its workers, which would time the
        # reference on one thread at N > 1
        host_cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.set_threads(host_cores)
        self.cores = O.ref().ref_num_threads() if self.kind == "reference" else O.oracle().oracle_num_threads()
        t0 = time.perf_counter()
        self.b = O.ref_bvh_build(scene_buf) if self.kind == "reference" else O.bvh_build(scene_buf)
        self.build_ms = (time.perf_counter() - t0) * 1e3
        self.n = self.b["n"]

    def set_threads(self, n):
        (self.O.ref().ref_set_num_threads if self.kind == "reference" else self.O.oracle().oracle_set_num_threads)(int(n))

    def frame(self, cam):
        """-> (rays traced, seconds)"""
        O, b, n, scene = self.O, self.b, self.n, self.scene
        t0 = time.perf_counter()
        c = O.ref_trace_primary(cam, b["nodes"], n, scene) if self.kind == "reference" else O.bvh_trace_primary(cam, b)
        t1 = time.perf_counter()
        sr = O.shadow_rays(cam, scene, c)
        srl = np.ascontiguousarray(sr[sr["idx"] != 0xFFFFFFFF])
        t2 = time.perf_counter()
        if self.kind == "reference":
            O.ref_trace_rays(srl, b["nodes"], n, scene); O.ref_shade(cam, c, scene)
        else:
            O.bvh_trace_rays(srl, b); O.shade(cam, scene, c)
        t3 = time.perf_counter()
        return len(c) + len(srl), (t1 - t0) + (t3 - t2)


def cpu_baseline(scene_buf, res, budget_s=12.0, single_thread_budget_s=4.0):
    from openclraytracer_b200 import frame
    ref = CpuReference(scene_buf)
    rays, secs, frames = 0, 0.0, 0
    while frames < 1 or (secs < budget_s and frames < 64):
        r, s = ref.frame(frame.orbit_camera(scene_buf, res[0], res[1], frames))
        rays += r; secs += s; frames += 1
    sample = f"{frames} frame(s) of {res[0]}x{res[1]} primary+shadow rays ({rays} rays) on the bench camera path"
    # the same on ONE thread ("timed at 1 thread and at all host cores"), a quarter-size frame, ~4 s
    ref.set_threads(1)
    r1, s1, f1, res1 = 0, 0.0, 0, (res[0] // 2, res[1] // 2)
    while f1 < 1 or (s1 < single_thread_budget_s and f1 < 16):
        r, s = ref.frame(frame.orbit_camera(scene_buf, res1[0], res1[1], f1))
        r1 += r; s1 += s; f1 += 1
    ref.set_threads(ref.cores)
    ref.single_thread = {"value": r1 / s1 / 1e6, "unit": UNIT, "sample": f"{f1} frame(s) of {res1[0]}x{res1[1]} ({r1} rays)"}
    return rays / secs / 1e6, ref, sample


def run_reference(args, workload_desc, scene_buf, n_tris):
    """--impl reference: the reference's own CPU implementation on the host cores, same metric/config; each step is a
    bounded sample (one 960x540 frame of the same camera path).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from openclraytracer_b200 import frame
    res = (960, 540)
    ref = CpuReference(scene_buf)
    rays, secs = 0, 0.0
    for i in range(args.warmup + args.steps):
        r, s = ref.frame(frame.orbit_camera(scene_buf, res[0], res[1], i))
        if i >= args.warmup:
            rays += r; secs += s
    value = rays / secs / 1e6
    sample = f"{args.steps} frame(s) of {res[0]}x{res[1]} primary+shadow rays ({rays} rays) on the bench camera path"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": DATA[args.workload],
            "config": {"workload": workload_desc, "resolution": "1920x1080 (bounded sample: %dx%d per step)" % res, "accel": "LBVH",
                       "triangles": n_tris},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "build_ms_per_Mtri": ref.build_ms / (n_tris / 1e6)}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
def time_trace_kernels(r, cam, repeats=5):
    """The traversal kernels of one frame, each timed alone with CUDA events on the launching stream:
    trace_primary_kernel (closest hit, camera entry point) and shadow_shade_kernel (occlusion queries + shading).
    Algorithmic bytes from the kernels' statistics mode: 128 B per wide-node visit + 48 B per triangle test, + 32 B contact
    written (primary) / 32 B contact read + 4 B pixel written (shadow)."""
    import torch
    a = r.accel
    n = r.resX * r.resY
    L1 = r.scene.num_lights() + 1
    stream = r.stream
    with torch.cuda.stream(stream):
        contacts = torch.empty(n * 32, dtype=torch.uint8, device="cuda")
        rgba = torch.empty(n, dtype=torch.int32, device="cuda")

        def timed(fn):
            ms = []
            for _ in range(repeats + 2):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); fn(); e1.record(stream); e1.synchronize()
                ms.append(e0.elapsed_time(e1))
            return float(np.median(ms[2:]))

        prim = lambda: a.generateContacts(cam, contacts.data_ptr())
        shad = lambda: a.shadow_shade(cam, contacts.data_ptr(), n, rgba.data_ptr())
        a.set_stats(True); prim(); sp = a.read_stats()
        a.set_stats(True); shad(); ss = a.read_stats()
        a.set_stats(False)
        ms_p, ms_s = timed(prim), timed(shad)
    bytes_p = sp[1] * 128 + sp[2] * 48 + n * 32            # 128-B wide node per visit, 48-B triangle per test, 32-B contact out
    bytes_s = ss[1] * 128 + ss[2] * 48 + n * 36
    return {"primary": {"ms": ms_p, "rays": sp[0], "bytes": bytes_p, "inner_per_ray": sp[1] / max(sp[0], 1), "tris_per_ray": sp[2] / max(sp[0], 1)},
            "shadow": {"ms": ms_s, "rays": ss[0], "bytes": bytes_s, "inner_per_ray": ss[1] / max(ss[0], 1), "tris_per_ray": ss[2] / max(ss[0], 1)}}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="soup1m", choices=["clumsydragon", "soup1m"],
                    help="soup1m = BASELINE.json configs[2] (the config the metric is quoted on at 1/2/4/8 GPUs; default), clumsydragon = configs[1]")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel build/sort/scan roofline section")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    scene_buf, workload_desc = make_scene(args.workload)
    n_tris = int(np.frombuffer(scene_buf[:96].tobytes(), dtype=np.uint64)[7])
    if args.impl == "reference":
        return run_reference(args, workload_desc, scene_buf, n_tris)

    import torch
    import torch.distributed as dist
    from openclraytracer_b200 import api, frame

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    hbm_peak, peak_src = load_peaks()

    r = frame.FrameRenderer(scene_buf, RES[0], RES[1], "BVH", rank, world, local_rank)
    # ---- build (once; timed separately, median of several) ------------------------------------------------
    r.accel.enable_stage_timing(True)
    build_ms, stages = [], []
    for i in range(8):
        r.build()
        build_ms.append(r.accel.stage_ms(api.STAGE_TOTAL))
        stages.append([r.accel.stage_ms(s) for s in (0, 1, 2, 3, api.STAGE_TRACE_TREE)])
    r.accel.enable_stage_timing(False)
    r.build()
    build_med = float(np.median(build_ms[2:]))
    stage_med = np.median(np.array(stages[2:]), axis=0)

    cams = [frame.orbit_camera(scene_buf, RES[0], RES[1], s) for s in range(args.warmup + args.steps)]
    stream = r.stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # rays per frame (counted once, untimed): primary = pixels of the stripe, shadow = live shadow rays
    rays_per_step = np.zeros(args.warmup + args.steps, dtype=np.int64)
    for s in range(args.warmup):
        rays_per_step[s] = r.render_device(cams[s], True, want_rays=True)
    for s in range(args.warmup, args.warmup + args.steps):
        rays_per_step[s] = r.render_device(cams[s], True, want_rays=True)
    # ---- timed: device-resident -------------------------------------------------------------------------
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for s in range(args.warmup, args.warmup + args.steps):
        r.render_device(cams[s], True)
    r.join()                                       # device-side join of the trace / gather streams: every frame is complete at ev1
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1)
    # ---- timed: e2e (host camera in, pinned host raster out, every step) ---------------------------------
    for s in range(args.warmup):
        r.render_host(cams[s], True)
    barrier()
    t_e2e0 = torch.cuda.Event(enable_timing=True); t_e2e1 = torch.cuda.Event(enable_timing=True)
    t_e2e0.record(stream)
    for s in range(args.warmup, args.warmup + args.steps):
        r.render_host_async(cams[s], True)         # two frames in flight; every frame's raster lands in pinned host memory
    r.join()
    t_e2e1.record(stream)
    barrier()
    e2e_ms = t_e2e0.elapsed_time(t_e2e1)
    clocks = sampler.stop() if rank == 0 else None

    # ---- dominant kernel, timed alone on its stream (the primary + shadow traversal launches of one frame) ----
    kt = time_trace_kernels(r, cams[args.warmup], repeats=5)

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    rays = torch.tensor([int(rays_per_step[args.warmup:].sum())], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    total_rays = int(rays[0])

    if rank == 0:
        value = total_rays / (dev_ms * 1e-3) / 1e6
        e2e_value = total_rays / (e2e_ms * 1e-3) / 1e6
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": DATA[args.workload],
            "config": {"workload": workload_desc, "resolution": "1920x1080", "accel": "LBVH", "triangles": n_tris,
                       "rays_per_step": total_rays // args.steps, "shadow": "1 per hit per light (scene light + camera light)",
                       "parallelism": f"image-space tile stripes x{world}, redundant build, " + EXCHANGE_DESC[r.exchange],
                       "l2": "per-step working set (contacts + shadow rays + shadow contacts, >400 MB at 1080p) exceeds the 126 MB L2; "
                             "the camera moves every step; the BVH stays L2 resident as in steady-state rendering"},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 64, "d2h_bytes_per_step": RES[0] * RES[1] * 4,
                    "ms_per_step": e2e_ms / args.steps},
            # trace_primary_kernel, shadow_shade_kernel + untile_kernel (nccl) | push, 2 signal, 2 wait (push) | ack wait, exchange (fused)
            "gpu_launches": args.steps * {"nccl": 3, "push": 7, "fused": 4}[r.exchange],
            "pipeline": f"{r.DEPTH} frames in flight: frames traced on alternating streams (frame slots)" +
                        (", gather + untile (+ read-back) on one more" if r.exchange == "nccl" else ", exchange + read-back on the frame's own stream"),
            "build_ms": build_med, "build_ms_per_Mtri": build_med / (n_tris / 1e6),
            "build_stage_ms": {"morton": float(stage_med[0]), "sort": float(stage_med[1]), "emit": float(stage_med[2]), "refit": float(stage_med[3]),
                               "trace_structure": float(stage_med[4])},
            "trace_structure": "separate trace tree + 4-wide collapse" if r.accel.has_trace_tree() else "4-wide collapse of the reference tree",
            "clocks": clocks,
        }
        # roofline of the dominant kernel (the longer of the two traversal kernels of the frame)
        dom = "primary" if kt["primary"]["ms"] >= kt["shadow"]["ms"] else "shadow"
        kd = kt[dom]
        achieved = kd["bytes"] / (kd["ms"] * 1e-3) / 1e9
        line["roofline"] = {"bound": "hbm", "kernel": "trace_primary_kernel" if dom == "primary" else "shadow_shade_kernel",
                            "achieved": achieved, "peak": hbm_peak, "unit": "GB/s", "frac": achieved / hbm_peak,
                            "traffic": ncu_traffic(args.workload, "trace_primary_kernel" if dom == "primary" else "shadow_shade_kernel"),
                            "peak_source": peak_src, "ms_per_launch": kd["ms"], "rays": kd["rays"],
                            "mrays_per_s_kernel_only": kd["rays"] / kd["ms"] / 1e3,
                            "note": "traversal is latency/divergence bound and served by L1/L2 (DESIGN.md section 4): the fraction explains, "
                                    "it does not bound; HBM-bound kernels are under 'kernels'"}
        line["trace_kernels"] = {k: {"ms": v["ms"], "rays": v["rays"], "mrays_per_s": v["rays"] / v["ms"] / 1e3,
                                     "inner_visits_per_ray": v["inner_per_ray"], "tri_tests_per_ray": v["tris_per_ray"],
                                     "algorithmic_GBs": v["bytes"] / (v["ms"] * 1e-3) / 1e9} for k, v in kt.items()}
        if not args.no_kernels:
            line["kernels"] = kernel_rooflines(r, hbm_peak, n_tris, stage_med)
        if not args.no_cpu_baseline and world == 1:          # the CPU baseline is reported at N = 1 only
            v, ref, sample = cpu_baseline(scene_buf, (960, 540))
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "sample": sample,
                                    "single_thread": ref.single_thread,
                                    "build_ms_per_Mtri": ref.build_ms / (n_tris / 1e6)}
        print(json.dumps(line))
    r.close()
    if world > 1:
        dist.destroy_process_group()


EXCHANGE_DESC = {"nccl": "all-gather of RGBA8 stripes",
                 "push": "stripes pushed into every GPU's framebuffer over peer memory (no collective)",
                 "fused": "pixels stored by the shading kernel into every GPU's framebuffer over peer memory (no collective)"}


def ncu_traffic(workload, kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of `kernel` from the committed `ncu --set full` capture of this
    workload (profiles/r01_traffic.json, written by profiles/summarize.py); None when no capture is committed."""
    try:
        t = json.load(open(os.path.join(ROOT, "profiles", "r01_traffic.json")))
        return t.get(f"{workload}_{kernel}")
    except Exception:
        return None


def kernel_rooflines(r, hbm_peak, n_tris, stage_med):
    """HBM-bound kernels timed live (CUDA events on the context's stream): build stages on the bench scene, and the
    Sort / Prefix Sum building blocks at 2^24 pairs / elements (SURVEY.md 8d algorithmic bytes)."""
    import torch
    from openclraytracer_b200 import api
    out = []
    per_tri = {"morton": 104, "sort": 16, "emit": 64, "refit": 100}
    for i, k in enumerate(["morton", "sort", "emit", "refit"]):
        ms = float(stage_med[i])
        gbs = per_tri[k] * n_tris / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
        out.append({"kernel": f"bvh.{k}", "units": n_tris, "bytes_per_unit": per_tri[k], "ms": ms, "achieved_GBs": gbs, "frac": gbs / hbm_peak})
    ctx = r.ctx
    stream = r.stream
    with torch.cuda.stream(stream):
        for logn in (24, 26):
            n = 1 << logn
            g = torch.Generator(device="cuda"); g.manual_seed(1)
            keys = torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device="cuda", generator=g)
            pairs = torch.stack([keys, torch.arange(n, dtype=torch.int32, device="cuda")], dim=1).contiguous()
            work = torch.empty_like(pairs)
            srt = api.BitonicSort(ctx)
            ms = []
            for it in range(6):
                work.copy_(pairs)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); srt.sort(work.data_ptr(), n); e1.record(stream); e1.synchronize()
                ms.append(e0.elapsed_time(e1))
            m = float(np.median(ms[2:]))
            gbs = 16.0 * n / (m * 1e-3) / 1e9
            out.append({"kernel": "sort_pairs_u32 (onesweep, 4x8 bit)", "units": n, "bytes_per_unit": 16, "ms": m, "achieved_GBs": gbs,
                        "frac": gbs / hbm_peak, "Mpairs_per_s": n / m / 1e3})
            vals = torch.randint(0, 8, (n,), dtype=torch.int32, device="cuda", generator=g)
            outb = torch.empty_like(vals)
            ps = api.PrefixSum(ctx)
            ms = []
            for it in range(6):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); ps.computePrefixSum(vals.data_ptr(), outb.data_ptr(), n); e1.record(stream); e1.synchronize()
                ms.append(e0.elapsed_time(e1))
            m = float(np.median(ms[2:]))
            gbs = 8.0 * n / (m * 1e-3) / 1e9
            out.append({"kernel": "scan_inclusive_u32 (decoupled look-back)", "units": n, "bytes_per_unit": 8, "ms": m, "achieved_GBs": gbs, "frac": gbs / hbm_peak})
            del keys, pairs, work, vals, outb
    return out


if __name__ == "__main__":
    main()
