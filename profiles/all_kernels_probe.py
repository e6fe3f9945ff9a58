"""profiles/all_kernels_probe.py — launches EVERY kernel of libclrt_cuda.so at least once at bench sizes (Soup-1M, 1080p, 2^24 blocks), so
that one `ncu --set full` pass over it yields a capture per kernel (north star: "every kernel with an ncu capture").  Few launches per
kernel: the pass replays each launch ~40 times.  usage: all_kernels_probe.py [soup triangles] (default 1 000 000)"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("CLRT_FRAME_DEPTH", "2")
import numpy as np, torch
from openclraytracer_b200 import api, frame, scenes, layouts as L

N = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
W, H = 1920, 1080
S = torch.cuda.Stream()
torch.cuda.set_stream(S)
ctx = api.Context(0, stream=S.cuda_stream)
buf = scenes.soup_scene(N)
scene = api.Scene(buf); scene.loadToGPU(ctx)                       # flatten_triangles_kernel
cam = frame.orbit_camera(buf, W, H, 5)
n = W * H

# ---- building blocks --------------------------------------------------------------------------------------------------------
e = 24
keys = torch.randint(0, 1 << 30, (1 << e,), dtype=torch.int32, device="cuda")
pairs = torch.stack([keys, torch.arange(1 << e, dtype=torch.int32, device="cuda")], dim=1).contiguous()
api.BitonicSort(ctx).sort(pairs.data_ptr(), 1 << e)              # radix_histogram, radix_scan_hist, onesweep_pass x4
vals = torch.randint(0, 8, (1 << e,), dtype=torch.int32, device="cuda"); out = torch.empty_like(vals)
api.PrefixSum(ctx).computePrefixSum(vals.data_ptr(), out.data_ptr(), 1 << e)      # scan_window_kernel<4>
small = pairs[: 1 << 12].clone()
api.BitonicSort(ctx, mode=api.SORT_BITONIC_COMPAT).sort(small.data_ptr(), 1 << 12)  # bitonic_stage_kernel
odd = torch.zeros((1 << 20) + 1, dtype=torch.int32, device="cuda")
api.PrefixSum(ctx).computePrefixSum(odd[1:].data_ptr(), odd[1:].data_ptr(), 1 << 20)    # scan_lookback_kernel (4-byte aligned only)
smallv = torch.ones(1 << 19, dtype=torch.int32, device="cuda")
api.PrefixSum(ctx).computePrefixSum(smallv.data_ptr(), smallv.data_ptr(), 1 << 19)    # scan_window_kernel<2>
del keys, pairs, vals, out, small, odd, smallv

# ---- LBVH: build (reference tree + separate trace tree), traversal, shading ----------------------------------------------------------
bvh = api.BVHManager(ctx, scene); bvh.initialize()
bvh.set_trace_tree(api.TRACE_TREE_ALWAYS); bvh.initializeFrame(); bvh.construct()    # + tt_morton, pyr_levels (trace-tree order), tt_emit
bvh.set_trace_tree(api.TRACE_TREE_AUTO); bvh.initializeFrame(); bvh.construct()      # morton_leaves, pyr_levels, karras_emit, wide_frame, widen, wide_top
contacts = torch.empty(n * 32, dtype=torch.uint8, device="cuda"); rgba = torch.empty(n, dtype=torch.int32, device="cuda")
L1 = scene.num_lights() + 1
rays = torch.empty(n * L1 * 48, dtype=torch.uint8, device="cuda"); scont = torch.empty(n * L1 * 32, dtype=torch.uint8, device="cuda")
bvh.set_trace_mode(api.TRACE_FAST)
bvh.generateContacts(cam, contacts.data_ptr())                                     # trace_primary_kernel<FAST>
bvh.shadow_shade(cam, contacts.data_ptr(), n, rgba.data_ptr())                     # shadow_shade_kernel
stripe = torch.empty(frame.stripe_capacity(W, H, 1), dtype=torch.int32, device="cuda")
bvh.render_stripe(cam, True, 0, 1, stripe.data_ptr())                              # trace_primary + shadow_shade (two-kernel frame)
bvh.set_fused_shadows(2)
bvh.render_stripe(cam, True, 0, 1, stripe.data_ptr())                              # frame_kernel (primary + occlusion + shading in one launch)
bvh.set_fused_shadows(1)
bvh.render_stripes_batch(S.cuda_stream, 0, [cam, frame.orbit_camera(buf, W, H, 6)], 0, 1, torch.empty(2 * frame.stripe_capacity(W, H, 1), dtype=torch.int32, device="cuda").data_ptr())   # trace_primary_batch, shadow_shade_batch
api.frame_untile(ctx, stripe.data_ptr(), W, H, 1, rgba.data_ptr())                 # untile_kernel
api.generate_shadow_rays(ctx, scene, cam, contacts.data_ptr(), n, rays.data_ptr())   # shadow_rays_kernel
bvh.generateContacts(rays.data_ptr(), scont.data_ptr(), n * L1)                    # trace_rays_kernel<FAST>
api.shade(ctx, scene, cam, contacts.data_ptr(), n, scont.data_ptr(), rgba.data_ptr(), None)    # shade_kernel
api.generate_primary_rays(ctx, cam, rays.data_ptr())                               # primary_rays_kernel
api.generate_bounce_rays(ctx, cam, contacts.data_ptr(), n, 20161116, rays.data_ptr())          # secondary_rays_kernel<DIFFUSE>
bvh.generateContacts(rays.data_ptr(), scont.data_ptr(), n)
api.generate_secondary_rays(ctx, api.RAYS_REFLECT, None, rays.data_ptr(), scont.data_ptr(), n, 0, rays.data_ptr())   # <REFLECT, FROM_RAYS>
api.shade_bounce(ctx, scene, cam, contacts.data_ptr(), n, scont.data_ptr(), rgba.data_ptr())   # shade_bounce_kernel
small_cam = frame.orbit_camera(buf, 480, 270, 5); small_cam["supersampilngFactor"] = 2
bvh.render_frame_host(small_cam, True)                                              # downsample_kernel
bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
qcam = frame.orbit_camera(buf, 480, 270, 5)
bvh.generateContacts(qcam, contacts.data_ptr())                                    # trace_primary_kernel<REFERENCE_ORDER> (quarter frame: unpruned)
api.generate_primary_rays(ctx, qcam, rays.data_ptr())
bvh.generateContacts(rays.data_ptr(), scont.data_ptr(), 480 * 270)                 # trace_rays_kernel<REFERENCE_ORDER>
bvh.set_trace_mode(api.TRACE_FAST)
ctx.finish()

# ---- Two-Level Grid and KD-tree ---------------------------------------------------------------------------------------------------
tlg = api.TwoLevelGridManager(ctx, scene); tlg.initialize(); tlg.initializeFrame(); tlg.construct()   # tlg_* build kernels, extract_ranges
tlg.generateContacts(cam, contacts.data_ptr())                                     # tlg_trace_primary_kernel<literal>
tlg.set_trace_mode(api.TRACE_FAST)
tlg.generateContacts(cam, contacts.data_ptr())                                     # tlg_walk_primary_kernel (corrected walk, lanes refilled)
api.generate_primary_rays(ctx, cam, rays.data_ptr())
tlg.generateContacts(rays.data_ptr(), scont.data_ptr(), n)                         # tlg_walk_rays_kernel
tlg.set_trace_mode(api.TRACE_REFERENCE_ORDER)
tlg.generateContacts(rays.data_ptr(), scont.data_ptr(), n)                         # tlg_trace_rays_kernel<literal>
kd = api.KDTreeManager(ctx, scene); kd.initialize(); kd.initializeFrame(); kd.construct()             # kd_* build kernels
kd.generateContacts(cam, contacts.data_ptr())                                      # kd_walk_primary_kernel
kd.generateContacts(rays.data_ptr(), scont.data_ptr(), n)                          # kd_walk_rays_kernel
ctx.finish()
for m in (kd, tlg, bvh):
    m.free()
del contacts, rgba, rays, scont, stripe

# ---- the frame exchanges over peer memory (one GPU: it is its own peer) --------------------------------------------------------------
for exchange in ("fused", "push"):
    os.environ["CLRT_EXCHANGE"] = exchange
    r = frame.FrameRenderer(buf, W, H, "BVH", 0, 1, 0)
    r.build()
    for s in range(3):
        r.render_host_async(frame.orbit_camera(buf, W, H, s), True)               # wait, frame_kernel + peer stores, signal / exchange, push_kernel
    r.finish(); r.close()
scene.free(); ctx.close()
print("all kernels launched")
