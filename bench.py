#!/usr/bin/env python
"""bench.py — the headline benchmark of BASELINE.json on N B200s of one box.

metric   : Mrays/s (primary + shadow rays, 1920x1080), whole job.  Also reported: BVH build ms per M triangles.
workload : default `soup1m` = BASELINE.json configs[2], the configuration the metric is quoted on at 1/2/4/8 B200 (synthetic
           1M-triangle random soup, 46 sub-meshes); `--accel GRID` walks the Two-Level Grid instead of the LBVH (configs[2]'s
           "LBVH vs Two-Level Grid").  `--workload clumsydragon` = configs[1], the largest bundled TestModels mesh present
           (50 032 triangles; Skull / swfighter are missing blobs).  Build once, then per step ONE 1080p frame: primary rays, one
           shadow ray per hit per light (scene light + the reference's camera light), Blinn-Phong shading.
           `--workload spheres10m` = configs[3]: 10M-triangle sphere field, 3840x2160, primary rays + one cosine-weighted diffuse
           bounce per hit through the generic-ray entry point, image-tile sharded (metric: Mrays/s primary+bounce at 4K).
step     : one frame of the moving demo camera.  N > 1: the frame is sharded in image space (tile stripes); every rank ends up
           with the whole frame — total work fixed => "scaling": "strong".
value    : rays traced by all ranks / device time (CUDA events on the launching stream, max over ranks), with the
           camera and the built structure already resident.
e2e      : the same through the renderer's public call (what a user calls; frames in flight): host camera in, pinned-host RGBA8 raster out.
roofline : the dominant kernel of the step is BVH traversal, which is bound by INSTRUCTION ISSUE at divergent warps, not by HBM or
           tensor throughput: achieved = algorithmic lane-ops/s (60 per binary inner-node visit = 120 per 4-wide
           visit, 75 per triangle test; visits and tests counted by the kernel's statistics mode) against the FP32 issue peak
           SMs x 128 lanes x SM clock.  `traffic` = DRAM bytes of that launch from the committed ncu capture.  The HBM-bound kernels
           (build stages, sort, scan: configs[4]'s 2^20..2^28 sweep) are reported per kernel under "kernels" against the measured
           copy bandwidth.
cpu_baseline / --impl reference : the reference's own CPU code for the path (oracle/_ref/libclrt_ref.so = its headers
           compiled in place; else the oracle port) on the host cores, frames of the SAME resolution and camera path.
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

UNIT = "Mrays/s"
WORKLOADS = {
    # name: (metric, resolution, data)
    "soup1m": ("Mrays/s (primary+shadow, 1080p)", (1920, 1080), "synthetic"),
    "clumsydragon": ("Mrays/s (primary+shadow, 1080p)", (1920, 1080), "bundled mesh, synthetic camera path"),
    "spheres10m": ("Mrays/s (primary+diffuse bounce, 4K)", (3840, 2160), "synthetic"),
}
ACCEL_NAME = {"BVH": "LBVH", "GRID": "Two-Level Grid (corrected closest-hit walk)", "KD": "KD-tree (ours; the reference's TODO 2.1)"}
LANE_OPS_PER_WIDE_VISIT = 120.0      # ~60 lane-ops per binary inner-node visit (two slab tests); a 4-wide visit = two of them
LANE_OPS_PER_TRI_TEST = 75.0         # Moeller-Trumbore incl. normal


def load_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return float(d["hbm_gbs"]), float(d.get("sm_max_mhz", 1965.0)), "measured (MEASURED_PEAKS.json)"
    return 6650.0, 1965.0, "fallback (B200_PROFILING.md)"


def make_scene(workload):
    from openclraytracer_b200 import scenes
    if workload == "clumsydragon":
        return scenes.load_bundled("clumsydragon"), "ClumsyDragon (50032 tris, bundled TestModels) + LIGHT 0 10 -330 1000"
    if workload == "soup1m":
        return scenes.soup_scene(1_000_000), "Soup-1M (1e6 random triangles, 46 sub-meshes, seed 20161116)"
    if workload == "spheres10m":
        return scenes.sphere_field_scene(), "Spheres-10M (2000 tessellated spheres x 5000 triangles, seed 20161116)"
    raise SystemExit(f"unknown workload {workload}")


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "50"],
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
    """The reference's CPU implementation of the step.  BVH frame: primary rays through the reference's own bvh_generate_contact,
    one shadow ray per hit per light through its generic-ray path (the reference has no shadow rays; they are generated by the
    oracle, untimed), cpuShadeBlinnPhong.  GRID frame: the same through the reference's TwoLevelGrid.h.  Bounce frame: primary rays +
    one diffuse bounce ray per hit (generated by the oracle-side numpy restatement, untimed) through the generic-ray path.
    oracle/_ref/libclrt_ref.so (the reference's headers compiled in place, OpenMP over work-items) when present ("reference"), else the
    oracle port ("port")."""

    def __init__(self, scene_buf, accel="BVH", force_port=False):
        from oracle import oracle_py as O
        O.build(("oracle",))
        self.O, self.scene, self.accel = O, scene_buf, accel
        # force_port: Spheres-10M has 2000 sub-meshes and the reference fetches a triangle by walking the sub-mesh headers
        # (SceneBufferParser.h:253-293, O(#sub-meshes) per fetch): its literal code needs minutes per 4K frame.  The oracle port reads the
        # flattened triangle array instead (validated against the literal walk on small scenes) — same arithmetic, bounded time.
        self.kind = "reference" if (O.have_ref() and not force_port) else "port"
        if accel == "KD":
            accel = self.accel = "BVH"                  # the reference has no KD-tree: its BVH is the CPU arm of that workload
        if accel == "GRID" and self.kind != "reference":
            raise SystemExit("the Two-Level Grid has no oracle port: oracle/_ref/libclrt_ref.so (the reference's own code) is required")
        # all the host cores this process may run on — torchrun exports OMP_NUM_THREADS=1 to its workers, which would time the
        # reference on one thread at N > 1
        host_cores = len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)
        self.set_threads(host_cores)
        self.cores = O.ref().ref_num_threads() if self.kind == "reference" else O.oracle().oracle_num_threads()
        t0 = time.perf_counter()
        if accel == "GRID":
            self.g = O.RefTlg(scene_buf)
            self.n = int(np.frombuffer(scene_buf[:96].tobytes(), dtype=np.uint64)[7])
        else:
            self.b = O.ref_bvh_build(scene_buf) if self.kind == "reference" else O.bvh_build(scene_buf)
            self.n = self.b["n"]
        self.build_ms = (time.perf_counter() - t0) * 1e3

    def set_threads(self, n):
        (self.O.ref().ref_set_num_threads if self.kind == "reference" else self.O.oracle().oracle_set_num_threads)(int(n))

    def _primary(self, cam):
        O = self.O
        if self.accel == "GRID":
            return self.g.trace_primary(cam)
        return O.ref_trace_primary(cam, self.b["nodes"], self.n, self.scene) if self.kind == "reference" else O.bvh_trace_primary(cam, self.b)

    def _rays(self, rays):
        O = self.O
        if self.accel == "GRID":
            return self.g.trace_rays(rays)
        return O.ref_trace_rays(rays, self.b["nodes"], self.n, self.scene) if self.kind == "reference" else O.bvh_trace_rays(rays, self.b)

    def frame(self, cam):
        """primary + shadow + shading -> (rays traced, seconds)"""
        O, scene = self.O, self.scene
        t0 = time.perf_counter()
        c = self._primary(cam)
        t1 = time.perf_counter()
        sr = O.shadow_rays(cam, scene, c)
        srl = np.ascontiguousarray(sr[sr["idx"] != 0xFFFFFFFF])
        t2 = time.perf_counter()
        self._rays(srl)
        if self.kind == "reference":
            O.ref_shade(cam, c, scene)
        else:
            O.shade(cam, scene, c)
        t3 = time.perf_counter()
        return len(c) + len(srl), (t1 - t0) + (t3 - t2)

    def bounce_frame(self, cam, seed=20161116):
        """primary + one diffuse bounce ray per hit + shading -> (rays traced, seconds); ray generation untimed"""
        O, scene = self.O, self.scene
        t0 = time.perf_counter()
        c = self._primary(cam)
        t1 = time.perf_counter()
        br = bounce_rays_host(O, cam, c, seed)
        t2 = time.perf_counter()
        self._rays(br)
        if self.kind == "reference":
            O.ref_shade(cam, c, scene)
        else:
            O.shade(cam, scene, c)
        t3 = time.perf_counter()
        return len(c) + len(br), (t1 - t0) + (t3 - t2)


def bounce_rays_host(O, cam, contacts, seed):
    """numpy restatement of secondary_rays_kernel<DIFFUSE> (csrc/shade.cu) for the live contacts -> Ray array (untimed helper)."""
    from openclraytracer_b200 import layouts as L, scenes
    hit = np.flatnonzero(contacts["nd"][:, 3] > 0)
    prim = O.generate_primary_rays(cam)
    o = prim["origin"][hit, :3].astype(np.float64); d = prim["direction"][hit, :3].astype(np.float64)
    t = contacts["nd"][hit, 3].astype(np.float64)[:, None]
    n = contacts["nd"][hit, :3].astype(np.float64).copy()
    n[(n * d).sum(axis=1) > 0] *= -1
    P = o + d * t
    with np.errstate(over="ignore"):
        u = scenes.splitmix64(np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + contacts["pixelIndex"][hit].astype(np.uint64))
    u1 = (u >> np.uint64(40)).astype(np.float64) / 16777216.0
    u2 = (u & np.uint64(0xFFFFFF)).astype(np.float64) / 16777216.0
    r, phi = np.sqrt(u1), 2 * np.pi * u2
    dx, dy, dz = r * np.cos(phi), r * np.sin(phi), np.sqrt(np.maximum(0.0, 1.0 - u1))
    ax = np.where((np.abs(n[:, 0]) < 0.5)[:, None], np.array([1.0, 0, 0]), np.array([0, 1.0, 0]))
    t1 = np.cross(ax, n); t1 /= np.linalg.norm(t1, axis=1, keepdims=True) + 1e-300
    t2 = np.cross(n, t1)
    dirs = t1 * dx[:, None] + t2 * dy[:, None] + n * dz[:, None]
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True) + 1e-300
    rays = np.zeros(len(hit), dtype=L.RAY)
    rays["idx"] = contacts["pixelIndex"][hit]
    rays["origin"][:, :3] = (P + n * (1e-4 * t)).astype(np.float32)
    rays["direction"][:, :3] = dirs.astype(np.float32)
    return rays


def cpu_frame_fn(ref, workload):
    return ref.bounce_frame if workload == "spheres10m" else ref.frame


def cpu_sample_resolution(workload, res):
    """The CPU arm's frame: the bench resolution, except Spheres-10M (a quarter of 4K per axis: a 4K frame is minutes of CPU work)."""
    return (res[0] // 4, res[1] // 4) if workload == "spheres10m" else res


def cpu_baseline(scene_buf, res, workload, accel, budget_s=14.0, single_thread_budget_s=4.0):
    """Bounded sample of the same workload on the host cores: whole frames of the bench camera path for ~budget_s."""
    from openclraytracer_b200 import frame
    ref = CpuReference(scene_buf, accel, force_port=(workload == "spheres10m"))
    res = cpu_sample_resolution(workload, res)
    fn = cpu_frame_fn(ref, workload)
    rays, secs, frames = 0, 0.0, 0
    while frames < 1 or (secs < budget_s and frames < 64):
        r, s = fn(frame.orbit_camera(scene_buf, res[0], res[1], frames))
        rays += r; secs += s; frames += 1
    sample = f"{frames} frame(s) of {res[0]}x{res[1]} ({rays} rays) on the bench camera path, {secs:.1f} s of CPU work"
    # the same on ONE thread ("timed at 1 thread and at all host cores"), a 1/16-size frame, ~4 s
    ref.set_threads(1)
    r1, s1, f1, res1 = 0, 0.0, 0, (res[0] // 4, res[1] // 4)
    while f1 < 1 or (s1 < single_thread_budget_s and f1 < 16):
        r, s = fn(frame.orbit_camera(scene_buf, res1[0], res1[1], f1))
        r1 += r; s1 += s; f1 += 1
    ref.set_threads(ref.cores)
    ref.single_thread = {"value": r1 / s1 / 1e6, "unit": UNIT, "sample": f"{f1} frame(s) of {res1[0]}x{res1[1]} ({r1} rays)"}
    return rays / secs / 1e6, ref, sample


def run_reference(args, workload_desc, scene_buf, n_tris):
    """--impl reference: the reference's own CPU implementation on the host cores, same metric / config / resolution; each step is
    one whole frame of the bench camera path (a bounded sample: the caller's --steps frames).  Rank 0 only."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    from openclraytracer_b200 import frame
    metric, res, data = WORKLOADS[args.workload]
    full_res = res
    res = cpu_sample_resolution(args.workload, res)
    ref = CpuReference(scene_buf, args.accel, force_port=(args.workload == "spheres10m"))
    fn = cpu_frame_fn(ref, args.workload)
    rays, secs = 0, 0.0
    for i in range(args.warmup + args.steps):
        r, s = fn(frame.orbit_camera(scene_buf, res[0], res[1], i))
        if i >= args.warmup:
            rays += r; secs += s
    value = rays / secs / 1e6
    sample = f"{args.steps} frame(s) of {res[0]}x{res[1]} ({rays} rays) on the bench camera path"
    line = {"impl": "reference", "metric": metric, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": secs / args.steps * 1e3, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": data,
            "config": {"workload": workload_desc, "resolution": "%dx%d" % full_res + ("" if res == full_res else " (bounded sample: %dx%d frames)" % res),
                       "accel": ACCEL_NAME[args.accel], "triangles": n_tris, "rays_per_step": rays // max(args.steps, 1)},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "build_ms_per_Mtri": ref.build_ms / (n_tris / 1e6)}
    print(json.dumps(line))


# ---------------------------------------------------------------------------------------------------------------
def timed_launches(stream, fn, repeats=5):
    import torch
    ms = []
    for _ in range(repeats + 2):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream); fn(); e1.record(stream); e1.synchronize()
        ms.append(e0.elapsed_time(e1))
    return float(np.median(ms[2:]))


def time_trace_kernels(r, cam, repeats=5):
    """The traversal kernels of one BVH frame, each timed alone with CUDA events on the launching stream: trace_primary_kernel
    (closest hit, camera entry point) and shadow_shade_kernel (occlusion queries + shading).  Visits / tests from the kernels'
    statistics mode."""
    import torch
    a = r.accel
    n = r.resX * r.resY
    stream = r.stream
    with torch.cuda.stream(stream):
        contacts = torch.empty(n * 32, dtype=torch.uint8, device="cuda")
        rgba = torch.empty(n, dtype=torch.int32, device="cuda")
        prim = lambda: a.generateContacts(cam, contacts.data_ptr())
        shad = lambda: a.shadow_shade(cam, contacts.data_ptr(), n, rgba.data_ptr())
        a.set_stats(True); prim(); sp = a.read_stats()
        a.set_stats(True); shad(); ss = a.read_stats()
        a.set_stats(False)
        ms_p, ms_s = timed_launches(stream, prim, repeats), timed_launches(stream, shad, repeats)
    out = {}
    for k, ms, st, kern in (("primary", ms_p, sp, "trace_primary_kernel"), ("shadow", ms_s, ss, "shadow_shade_kernel")):
        out[k] = {"kernel": kern, "ms": ms, "rays": int(st[0]), "inner_per_ray": st[1] / max(st[0], 1), "tris_per_ray": st[2] / max(st[0], 1),
                  "lane_ops": st[1] * LANE_OPS_PER_WIDE_VISIT + st[2] * LANE_OPS_PER_TRI_TEST}
    return out


def issue_roofline(kd, workload, sm_count, sm_mhz, peak_src):
    """traversal is bound by instruction issue; fraction = algorithmic lane-ops/s over SMs x 128 lanes x clock."""
    peak = sm_count * 128 * sm_mhz * 1e6 / 1e12
    achieved = kd["lane_ops"] / (kd["ms"] * 1e-3) / 1e12
    prof = ncu_facts(workload, kd["kernel"])
    return {"bound": "issue", "kernel": kd["kernel"], "achieved": achieved, "peak": peak, "unit": "Tlane-op/s", "frac": achieved / peak,
            "traffic": prof.get("dram_bytes"), "lanes_per_instruction": prof.get("threads_per_inst"), "issue_active_pct": prof.get("issue_active_pct"),
            "peak_source": f"{sm_count} SMs x 128 FP32 lanes x {sm_mhz:.0f} MHz ({peak_src})", "ms_per_launch": kd["ms"], "rays": kd["rays"],
            "mrays_per_s_kernel_only": kd["rays"] / kd["ms"] / 1e3,
            "algorithmic_lane_ops_per_ray": kd["lane_ops"] / max(kd["rays"], 1),
            "note": "algorithmic lane-ops: 120 per 4-wide node visit (= two binary visits of 60), 75 per triangle test, counted by the kernel's "
                    "statistics mode; traversal is served by L1/L2 (traffic = DRAM bytes of one launch, ncu) — neither HBM nor tensor bound"}


def ncu_facts(workload, kernel):
    """Per-launch facts of `kernel` from the committed `ncu --set full` capture of this workload (profiles/r02_traffic.json, written by
    profiles/summarize.py): DRAM bytes, lanes per instruction, issue-active %; {} when no capture is committed."""
    for tag in ("r02", "r01"):
        try:
            t = json.load(open(os.path.join(ROOT, "profiles", f"{tag}_traffic.json")))
        except Exception:
            continue
        v = t.get(f"{workload}_{kernel}")
        if isinstance(v, dict):
            return v
        if v is not None:
            return {"dram_bytes": v}
    return {}


EXCHANGE_DESC = {"nccl": "all-gather of RGBA8 stripes",
                 "push": "stripes pushed into every GPU's framebuffer over peer memory (no collective)",
                 "fused": "pixels stored by the shading kernel into every GPU's framebuffer over peer memory (no collective)"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="soup1m", choices=sorted(WORKLOADS),
                    help="soup1m = BASELINE.json configs[2] (the config the metric is quoted on at 1/2/4/8 GPUs; default), clumsydragon = configs[1], "
                         "spheres10m = configs[3] (4K, primary + diffuse bounce)")
    ap.add_argument("--accel", default="BVH", choices=["BVH", "GRID", "KD"],
                    help="acceleration structure (-accStruct of the reference's demo; KD = the KD-tree manager, the reference's TODO 2.1)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-kernels", action="store_true", help="skip the per-kernel build/sort/scan roofline section")
    args = ap.parse_args()
    if args.impl == "reference" and args.steps == 200 and args.warmup == 10:
        args.steps, args.warmup = 20, 2                 # a CPU frame takes about a second
    if args.workload == "spheres10m" and args.accel != "BVH":
        raise SystemExit("spheres10m runs on the LBVH")
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
    hbm_peak, sm_mhz, peak_src = load_peaks()
    metric, RES, data = WORKLOADS[args.workload]
    bounce = args.workload == "spheres10m"

    if bounce:
        r = frame.BounceRenderer(scene_buf, RES[0], RES[1], rank, world, local_rank)
    else:
        r = frame.FrameRenderer(scene_buf, RES[0], RES[1], args.accel, rank, world, local_rank)
    sm_count = r.ctx.getMaxComputeUnits()
    # ---- build (once; timed separately, median of several) ------------------------------------------------
    r.accel.enable_stage_timing(True)
    build_ms, stages = [], []
    stage_ids = (0, 1, 2, 3, api.STAGE_TRACE_TREE) if args.accel == "BVH" else (8, 9, 10, 11, 12) if args.accel == "GRID" else (api.STAGE_TOTAL,)
    for i in range(8):
        r.build()
        build_ms.append(r.accel.stage_ms(api.STAGE_TOTAL))
        stages.append([r.accel.stage_ms(s) for s in stage_ids])
    r.accel.enable_stage_timing(False)
    r.build()
    build_med = float(np.median(build_ms[2:]))
    stage_med = np.median(np.array(stages[2:]), axis=0)

    total_steps = args.warmup + args.steps
    cams = [frame.orbit_camera(scene_buf, RES[0], RES[1], s) for s in range(total_steps)]
    stream = r.stream

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def submit(s, host):
        if bounce:
            (r.render_host if host else r.render_device)(cams[s])
        elif host:
            r.render_host_async(cams[s], True)
        else:
            r.render_device(cams[s], True)

    def join():
        r.join()

    # rays per frame (counted once, untimed): primary = pixels of the stripe, shadow / bounce = live secondary rays.  The camera path
    # has a period of 64 steps: count one period, reuse it.
    period = min(total_steps, 64)
    rays_per_pose = np.zeros(period, dtype=np.int64)
    for s in range(period):
        if bounce:
            r.render_device(cams[s]); rays_per_pose[s] = r.count_rays()
        else:
            rays_per_pose[s] = r.render_device(cams[s], True, want_rays=True)
    rays_timed = int(sum(rays_per_pose[s % 64] for s in range(args.warmup, total_steps)))
    # ---- timed: device-resident -------------------------------------------------------------------------
    for s in range(args.warmup):
        submit(s, False)
    sampler = ClockSampler(local_rank)
    barrier()
    if rank == 0:
        sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for s in range(args.warmup, total_steps):
        submit(s, False)
    join()                                         # device-side join of the trace / gather streams: every frame is complete at ev1
    ev1.record(stream)
    barrier()
    dev_ms = ev0.elapsed_time(ev1)
    # ---- timed: e2e (host camera in, pinned host raster out, every step) ---------------------------------
    for s in range(args.warmup):
        submit(s, True)
    r.finish()
    barrier()
    t_e2e0 = torch.cuda.Event(enable_timing=True); t_e2e1 = torch.cuda.Event(enable_timing=True)
    t_e2e0.record(stream)
    for s in range(args.warmup, total_steps):
        submit(s, True)                            # frames in flight; every frame's raster lands in pinned host memory
    join()
    t_e2e1.record(stream)
    barrier()
    e2e_ms = t_e2e0.elapsed_time(t_e2e1)
    clocks = sampler.stop() if rank == 0 else None
    r.accel.check()                                # a traversal-stack overflow anywhere above fails the run

    # ---- dominant kernel, timed alone on its stream ----------------------------------------------------------
    kt = None
    if rank == 0 and args.accel == "BVH":
        kt = time_bounce_kernels(r, cams[args.warmup]) if bounce else time_trace_kernels(r, cams[args.warmup], repeats=5)

    t = torch.tensor([dev_ms, e2e_ms], dtype=torch.float64, device="cuda")
    rays = torch.tensor([rays_timed], dtype=torch.int64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(rays, op=dist.ReduceOp.SUM)
    dev_ms, e2e_ms = float(t[0]), float(t[1])
    total_rays = int(rays[0])

    if rank == 0:
        value = total_rays / (dev_ms * 1e-3) / 1e6
        e2e_value = total_rays / (e2e_ms * 1e-3) / 1e6
        exchange = "nccl" if bounce else r.exchange
        if bounce:
            launches, pipeline = 5, f"{r.DEPTH} frames in flight (frame slots: own stream, counter and buffers each)"   # primary, bounce rays, trace, shade, untile (+ NCCL's own kernel at N > 1)
            secondary = "one cosine-weighted diffuse bounce ray per hit through generateContacts(rays, contacts, n)"
        else:
            # BVH: trace_primary_kernel, shadow_shade_kernel (or frame_kernel alone) + untile_kernel (nccl) | push, 2 signal, 2 wait (push) |
            # ack wait, ready signal, ready wait, ack signal (fused, split exchange)
            # GRID: primary, shadow-ray generation, generic trace, shade, untile
            launches = ({"nccl": 3, "push": 7, "fused": 5}[exchange] - (1 if r.one_kernel else 0)) if args.accel == "BVH" else 5
            if args.accel == "BVH" and exchange == "fused" and r.batch > 1:
                launches = 3 + 2.0 / r.batch                   # `batch` frames share one primary and one shadow + shade launch
            pipeline = f"{r.DEPTH} frame(s) in flight" + (": frames traced on alternating streams (frame slots)" if r.DEPTH > 1 else "") + \
                       (f", {r.batch} frames per pair of launches" if getattr(r, "batch", 1) > 1 else "")
            secondary = "1 shadow ray per hit per light (scene light + camera light)"
        line = {
            "metric": metric, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f32", "data": data,
            "config": {"workload": workload_desc, "resolution": "%dx%d" % RES, "accel": ACCEL_NAME[args.accel],
                       "triangles": n_tris, "rays_per_step": total_rays // args.steps, "secondary": secondary,
                       "parallelism": f"image-space tile stripes x{world}, redundant build, " + EXCHANGE_DESC[exchange],
                       "l2": "inputs larger than L2: the per-step working set (contacts, rays, framebuffers: > 130 MB at 1080p with frames in "
                             "flight) exceeds the 126 MB L2 and the camera moves every step; the acceleration structure stays cache resident "
                             "as in steady-state rendering",
                       "timed_region_s": dev_ms * 1e-3},
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": 64, "d2h_bytes_per_step": RES[0] * RES[1] * 4,
                    "ms_per_step": e2e_ms / args.steps},
            "gpu_launches": int(round(args.steps * launches)),
            "pipeline": pipeline,
            "build_ms": build_med, "build_ms_per_Mtri": build_med / (n_tris / 1e6),
            "build_stage_ms": dict(zip(["morton", "sort", "emit", "box_pyramid", "trace_structure"] if args.accel == "BVH" else
                                       ["count_top", "top_pairs", "leaf_cells", "leaf_count", "leaf_pairs"] if args.accel == "GRID" else ["total"],
                                       [float(x) for x in stage_med])),
            "clocks": clocks,
        }
        if args.accel == "BVH":
            line["trace_structure"] = "separate trace tree + 4-wide collapse" if r.accel.has_trace_tree() else "4-wide collapse of the reference tree"
        if kt:
            # roofline of the dominant kernel (the longest traversal kernel of the frame)
            dom = max(kt, key=lambda k: kt[k]["ms"])
            clock = (clocks or {}).get("sm_max_mhz") or sm_mhz
            line["roofline"] = issue_roofline(kt[dom], args.workload, sm_count, float(clock), peak_src)
            line["trace_kernels"] = {k: {"kernel": v["kernel"], "ms": v["ms"], "rays": v["rays"], "mrays_per_s": v["rays"] / v["ms"] / 1e3,
                                         "wide_visits_per_ray": v["inner_per_ray"], "tri_tests_per_ray": v["tris_per_ray"],
                                         "issue_frac": issue_roofline(v, args.workload, sm_count, float(clock), peak_src)["frac"]} for k, v in kt.items()}
        else:
            line["roofline"] = {"bound": "issue", "kernel": "tlg_trace_*_kernel" if args.accel == "GRID" else "kd_trace_*_kernel", "achieved": None, "peak": None,
                                "unit": "Tlane-op/s", "frac": None, "traffic": None,
                                "note": "this walk has no statistics mode (cells / nodes stepped per ray are not counted): see 'value'"}
        if not args.no_kernels:
            line["kernels"] = kernel_rooflines(r, hbm_peak, n_tris, stage_med, args.accel)
        if not args.no_cpu_baseline and world == 1:          # the CPU baseline is reported at N = 1 only
            v, ref, sample = cpu_baseline(scene_buf, RES, args.workload, args.accel)
            line["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": ref.cores, "kind": ref.kind, "sample": sample,
                                    "single_thread": ref.single_thread,
                                    "build_ms_per_Mtri": ref.build_ms / (n_tris / 1e6)}
        print(json.dumps(line))
    r.close()
    if world > 1:
        dist.destroy_process_group()


def time_bounce_kernels(r, cam, repeats=5):
    """configs[3]: the two traversal launches of a bounce frame, each timed alone: the stripe's primary rays (trace_primary_kernel) and
    the incoherent bounce rays through the generic-ray entry (trace_rays_kernel)."""
    import torch
    from openclraytracer_b200 import api
    a = r.accel
    with torch.cuda.stream(r.stream):
        prim = lambda: a.trace_primary_stripe(cam, r.rank, r.world, r.contacts.data_ptr())
        prim()
        api.generate_bounce_rays(r.ctx, cam, r.contacts.data_ptr(), r.slots, 20161116, r.rays.data_ptr())
        bnc = lambda: a.generateContacts(r.rays.data_ptr(), r.bounce.data_ptr(), r.slots)
        a.set_stats(True); prim(); sp = a.read_stats()
        a.set_stats(True); bnc(); sb = a.read_stats()
        a.set_stats(False)
        ms_p, ms_b = timed_launches(r.stream, prim, repeats), timed_launches(r.stream, bnc, repeats)
    out = {}
    for k, ms, st, kern in (("primary", ms_p, sp, "trace_primary_kernel"), ("bounce", ms_b, sb, "trace_rays_kernel")):
        out[k] = {"kernel": kern, "ms": ms, "rays": int(st[0]), "inner_per_ray": st[1] / max(st[0], 1), "tris_per_ray": st[2] / max(st[0], 1),
                  "lane_ops": st[1] * LANE_OPS_PER_WIDE_VISIT + st[2] * LANE_OPS_PER_TRI_TEST}
    return out


def kernel_rooflines(r, hbm_peak, n_tris, stage_med, accel):
    """HBM-bound kernels timed live (CUDA events on the context's stream): the build stages on the bench scene, and the Sort / Prefix
    Sum building blocks over BASELINE.json configs[4]'s sweep — 2^20 .. 2^28 (key, value) pairs / elements, uniform 30-bit keys and the
    mesh-like distribution with 50 % duplicate keys (SURVEY.md 8d algorithmic bytes: 16 B per pair, 8 B per element).  A 512 MB buffer
    is written between timed iterations (L2 flush): the small sizes would otherwise be served from the 126 MB L2."""
    import torch
    from openclraytracer_b200 import api
    out = []
    if accel == "BVH":
        # emit 64 B + refit 100 B per inner node are the ALGORITHMIC bytes of topology + boxes; this library does both in the
        # emit kernel from a box pyramid (csrc/boxpyr.cuh), so the two stages are one roofline entry (stage 3 = the pyramid, stage 2 = emit)
        for name, per, ms in (("bvh.morton", 104, float(stage_med[0])), ("bvh.sort", 16, float(stage_med[1])),
                              ("bvh.emit+boxes (box pyramid + emit)", 164, float(stage_med[2] + stage_med[3]))):
            gbs = per * n_tris / (ms * 1e-3) / 1e9 if ms > 0 else 0.0
            out.append({"kernel": name, "units": n_tris, "bytes_per_unit": per, "ms": ms, "achieved_GBs": gbs, "frac": gbs / hbm_peak})
        tot = float(sum(stage_med[:4]))
        if tot > 0:
            gbs = 284.0 * n_tris / (tot * 1e-3) / 1e9
            out.append({"kernel": "bvh.build (morton+sort+pyramid+emit)", "units": n_tris, "bytes_per_unit": 284, "ms": tot, "achieved_GBs": gbs, "frac": gbs / hbm_peak})
    ctx = r.ctx
    stream = r.stream
    free_b, _ = torch.cuda.mem_get_info()
    with torch.cuda.stream(stream):
        flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")
        srt, ps = api.BitonicSort(ctx), api.PrefixSum(ctx)

        def timed(fn, prep=None, reps=5):
            ms = []
            for it in range(reps + 1):
                if prep:
                    prep()
                flush.zero_()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(stream); fn(); e1.record(stream); e1.synchronize()
                ms.append(e0.elapsed_time(e1))
            return float(np.median(ms[1:]))

        for logn in (20, 22, 24, 26, 28):
            n = 1 << logn
            if n * 8 * 4 > free_b:
                continue
            g = torch.Generator(device="cuda"); g.manual_seed(1)
            idx = torch.arange(n, dtype=torch.int32, device="cuda")
            for dist_name in ("uniform", "50% duplicates"):
                keys = torch.randint(0, 1 << 30, (n,), dtype=torch.int32, device="cuda", generator=g)
                if dist_name != "uniform":
                    keys[1::2] = keys[0::2]                     # every key occurs (at least) twice: the mesh-like distribution
                pairs = torch.stack([keys, idx], dim=1).contiguous()
                work = torch.empty_like(pairs)
                m = timed(lambda: srt.sort(work.data_ptr(), n), prep=lambda: work.copy_(pairs))
                gbs = 16.0 * n / (m * 1e-3) / 1e9
                out.append({"kernel": f"sort_pairs_u32 2^{logn} {dist_name}", "units": n, "bytes_per_unit": 16, "ms": m, "achieved_GBs": gbs,
                            "frac": gbs / hbm_peak, "Gpairs_per_s": n / m / 1e6})
                del keys, pairs, work
            vals = torch.randint(0, 8, (n,), dtype=torch.int32, device="cuda", generator=g)
            outb = torch.empty_like(vals)
            m = timed(lambda: ps.computePrefixSum(vals.data_ptr(), outb.data_ptr(), n))
            gbs = 8.0 * n / (m * 1e-3) / 1e9
            out.append({"kernel": f"scan_inclusive_u32 2^{logn}", "units": n, "bytes_per_unit": 8, "ms": m, "achieved_GBs": gbs, "frac": gbs / hbm_peak})
            del vals, outb, idx
        del flush
    return out


if __name__ == "__main__":
    main()
