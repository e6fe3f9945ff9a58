"""GPU parity: scene flattening, LBVH build (Morton, sort, box pyramid, Karras emit + boxes), traversal, shading — through the C ABI,
against the committed golden vectors (reference's own code) and the oracle.  Integer/index work bit-exact; hit t /
normals bit-exact in REFERENCE_ORDER mode, FAST mode by the north-star tolerances."""
import numpy as np
import pytest
from conftest import BUNDLED, masked_nodes, digest
from openclraytracer_b200 import api, scenes, layouts as L

pytestmark = pytest.mark.gpu


def build_gpu(ctx, scene_buf):
    scene = api.Scene(scene_buf); scene.loadToGPU(ctx)
    bvh = api.BVHManager(ctx, scene); bvh.initialize(); bvh.initializeFrame(); bvh.construct()
    return scene, bvh


def check_hierarchy(nodes, n):
    """Testing/BVHTest.h:52-113 vectorised: every parent chain ends at the root; parent boxes contain child boxes."""
    parent = nodes["data"][:, 0].astype(np.int64)
    assert parent[n] == 0xFFFFFFFF
    cur = parent[:n].copy()
    for _ in range(4 * 64):
        live = cur != 0xFFFFFFFF
        if not live.any():
            break
        cur[live] = parent[cur[live]]
    assert (cur == 0xFFFFFFFF).all(), "a leaf's parent chain does not end at the root"
    inner = np.arange(n, 2 * n - 1)
    for c in (1, 2):
        child = nodes["data"][inner, c]
        assert (nodes["data"][child, 0] == inner).all()
        assert (nodes["bounds"][inner, 0, :3] <= nodes["bounds"][child, 0, :3]).all()
        assert (nodes["bounds"][inner, 1, :3] >= nodes["bounds"][child, 1, :3]).all()
    kids = np.concatenate([nodes["data"][inner, 1], nodes["data"][inner, 2]])
    assert len(np.unique(kids)) == 2 * n - 2 and n not in kids


@pytest.mark.parametrize("name", BUNDLED)
def test_build_matches_golden_and_oracle(ctx, O, golden, name):
    g = golden(name)
    buf = scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    n = scene.triangle_count()
    assert n == int(g["n"])
    tris, tim = O.flatten(buf)
    assert np.array_equal(scene.export_triangles(), tris)                 # flattening (incl. mis-aligned blocks, Q14)
    pairs, nodes = bvh.export_pairs(), bvh.export_nodes()
    assert digest(pairs) == str(g["pairs_sha"])                            # Morton codes + sorted order, bit-exact
    assert digest(masked_nodes(nodes, n)) == str(g["nodes_sha"])           # topology + boxes, bit-exact
    check_hierarchy(nodes, n)
    bvh.free(); scene.free()


@pytest.mark.parametrize("name", BUNDLED)
def test_trace_reference_order_bit_exact(ctx, golden, name):
    g = golden(name)
    buf = scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    for k in range(2):
        want = g[f"contacts{k}"]
        bvh.generateContacts(g[f"cam{k}"])
        got = bvh.download_primary_contacts()
        hit = want["nd"][:, 3] > 0
        assert np.array_equal(got["nd"][:, 3] > 0, hit)
        assert np.array_equal(got["nd"][hit], want["nd"][hit])
        assert np.array_equal(got["materialIndex"][hit], want["materialIndex"][hit])
        assert np.array_equal(got["pixelIndex"], want["pixelIndex"])
        assert (got["nd"][~hit] == 0).all()
    bvh.free(); scene.free()


def compare_fast(got, want, tris):
    """North-star tolerances: ids agree on >= 99.99 % of rays, mismatches only on near ties (|dt| <= 1e-5 t);
    t within 1e-4 relative."""
    hit = want["nd"][:, 3] > 0
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    tw, tg = want["nd"][hit, 3].astype(np.float64), got["nd"][hit, 3].astype(np.float64)
    assert (np.abs(tw - tg) <= 1e-4 * tw).all()
    same = got["pad"][hit, 0] == want["pad"][hit, 0]
    if not same.all():
        assert (np.abs(tw - tg)[~same] <= 1e-5 * tw[~same]).all()
    assert same.mean() >= 0.9999 or (~same).sum() <= 1
    nw, ng = want["nd"][hit][same][:, :3], got["nd"][hit][same][:, :3]
    assert np.array_equal(nw, ng)


@pytest.mark.parametrize("name", BUNDLED)
def test_trace_fast_within_tolerance(ctx, O, golden, name):
    g = golden(name)
    buf = scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    for cam in (g["cam0"], g["cam1"], api.demo_camera(buf, 320, 200)):
        want = O.bvh_trace_primary(cam, b)
        bvh.set_trace_mode(api.TRACE_FAST)
        bvh.generateContacts(cam)
        compare_fast(bvh.download_primary_contacts(), want, b["tris"])
    bvh.free(); scene.free()


@pytest.mark.parametrize("n,tpm", [(1, 24), (2, 24), (3, 24), (100, 16), (5000, 1000), (200_000, 21840)])
def test_soup_build_and_trace_vs_oracle(ctx, O, n, tpm):
    buf = scenes.soup_scene(n, seed=n, tris_per_mesh=tpm)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    assert np.array_equal(scene.export_triangles(), b["tris"])
    assert np.array_equal(bvh.export_pairs(), b["pairs"])
    nodes = bvh.export_nodes()
    assert np.array_equal(masked_nodes(nodes, n), masked_nodes(b["nodes"], n)[: 2 * n - 1])
    if n >= 2:
        check_hierarchy(nodes, n)
    cam = api.demo_camera(buf, 256, 144)
    want = O.bvh_trace_primary(cam, b)
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    got = bvh.download_primary_contacts()
    hit = want["nd"][:, 3] > 0
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    assert np.array_equal(got["nd"][hit], want["nd"][hit]) and np.array_equal(got["pad"][hit, 0], want["pad"][hit, 0])
    bvh.set_trace_mode(api.TRACE_FAST)
    bvh.generateContacts(cam)
    compare_fast(bvh.download_primary_contacts(), want, b["tris"])
    bvh.free(); scene.free()


def test_generic_rays_entry_point(ctx, O):
    """generateContacts(rays, contacts, rayCount): primary rays generated on the device == oracle's; traced contacts
    equal the camera entry point's; a zero-direction ray is a miss."""
    buf = scenes.load_bundled("gardenlight")
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    cam = api.demo_camera(buf, 128, 96)
    n = 128 * 96
    rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_primary_rays(ctx, cam, rays.getCLMem())
    r = rays.copyToHost(L.RAY, n)
    wr = O.generate_primary_rays(cam)
    assert np.array_equal(r["origin"][:, :3], wr["origin"][:, :3]) and np.array_equal(r["direction"][:, :3], wr["direction"][:, :3])
    r["direction"][5] = 0
    rays.upload(r)
    for mode in (api.TRACE_REFERENCE_ORDER, api.TRACE_FAST):
        bvh.set_trace_mode(mode)
        bvh.generateContacts(rays, contacts, n)
        got = contacts.copyToHost(L.CONTACT, n)
        want = O.bvh_trace_rays(r, b)
        assert got["nd"][5, 3] == 0
        if mode == api.TRACE_REFERENCE_ORDER:
            hit = want["nd"][:, 3] > 0
            assert np.array_equal(got["nd"][:, 3] > 0, hit) and np.array_equal(got["nd"][hit], want["nd"][hit])
        else:
            compare_fast(got, want, b["tris"])
    rays.free(); contacts.free(); bvh.free(); scene.free()


@pytest.mark.parametrize("name", ["cube", "tenryuu", "clumsydragon"])
def test_shading_matches_reference_host_loop(ctx, O, golden, name):
    """cpuShadeBlinnPhong (host loop in the reference) as a CUDA kernel: float colour within 1e-4, RGBA8 within 1 LSB."""
    g = golden(name)
    buf = scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    for k in range(2):
        cam = g[f"cam{k}"]
        bvh.generateContacts(cam)
        n = 96 * 64
        f32 = api.Buffer(ctx, n * 16); u8 = api.Buffer(ctx, n * 4)
        api.shade(ctx, scene, cam, bvh.getPrimaryContacts(), n, None, u8.getCLMem(), f32.getCLMem())
        got = f32.copyToHost(np.float32, n * 4).reshape(n, 4)
        want = g[f"colors{k}"]
        assert np.allclose(got, want, rtol=1e-4, atol=1e-5)
        want8 = np.floor(np.clip(want, 0, 1) * 255.0 + 0.5).astype(np.int32)
        got8 = u8.copyToHost(np.uint8, n * 4).reshape(n, 4).astype(np.int32)
        assert np.abs(got8 - want8).max() <= 1
        f32.free(); u8.free()
    bvh.free(); scene.free()


def test_shadow_pass_matches_oracle(ctx, O):
    buf = scenes.load_bundled("tenryuu")
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    cam = api.demo_camera(buf, 160, 100)
    n = 160 * 100
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    contacts = bvh.download_primary_contacts()
    L1 = scene.num_lights() + 1
    srays = api.Buffer(ctx, n * L1 * 48); scont = api.Buffer(ctx, n * L1 * 32); u8 = api.Buffer(ctx, n * 4)
    assert api.generate_shadow_rays(ctx, scene, cam, bvh.getPrimaryContacts(), n, srays.getCLMem()) == L1
    want_rays = O.shadow_rays(cam, buf, contacts)
    got_rays = srays.copyToHost(L.RAY, n * L1)
    live = want_rays["idx"] != 0xFFFFFFFF
    assert np.array_equal(got_rays["idx"] != 0xFFFFFFFF, live)
    assert np.allclose(got_rays["origin"][live], want_rays["origin"][live], rtol=1e-5, atol=1e-6)
    assert np.allclose(got_rays["direction"][live], want_rays["direction"][live], rtol=1e-5, atol=1e-6)
    bvh.generateContacts(srays, scont, n * L1)
    sc = scont.copyToHost(L.CONTACT, n * L1)
    want_sc = O.bvh_trace_rays(got_rays, b)
    assert np.array_equal(sc["nd"][:, 3], want_sc["nd"][:, 3])
    api.shade(ctx, scene, cam, bvh.getPrimaryContacts(), n, scont.getCLMem(), u8.getCLMem(), None)
    want8 = O.shade(cam, buf, contacts, shadow=sc).view(np.uint8).reshape(n, 4).astype(np.int32)
    got8 = u8.copyToHost(np.uint8, n * 4).reshape(n, 4).astype(np.int32)
    assert np.abs(got8 - want8).max() <= 1
    assert (sc["nd"][:, 3] > 0).any()                       # some light is actually occluded
    srays.free(); scont.free(); u8.free(); bvh.free(); scene.free()


@pytest.mark.parametrize("res", [(160, 100), (97, 61)])
def test_stripes_compose_to_the_single_gpu_frame(ctx, res):
    """Image-space sharding: stripes of nranks = 2, 3, 8 untiled == the one-rank frame (ragged edges included)."""
    buf = scenes.load_bundled("gardenlight")
    scene, bvh = build_gpu(ctx, buf)
    cam = api.demo_camera(buf, *res)
    ref_img, rays = bvh.render_frame_host(cam, shadows=True, want_rays=True)
    assert rays >= res[0] * res[1]
    for nranks in (2, 3, 8):
        cap, _ = api.frame_stripe_pixels(res[0], res[1], 0, nranks)
        stripes = api.Buffer(ctx, cap * 4 * nranks); raster = api.Buffer(ctx, res[0] * res[1] * 4)
        total = 0
        for r in range(nranks):
            total += bvh.render_stripe(cam, True, r, nranks, stripes.getCLMem() + r * cap * 4, want_rays=True)
        api.frame_untile(ctx, stripes.getCLMem(), res[0], res[1], nranks, raster.getCLMem())
        img = raster.copyToHost(np.uint8).reshape(res[1], res[0], 4)
        assert np.array_equal(img, ref_img)
        assert total == rays
        stripes.free(); raster.free()
    bvh.free(); scene.free()


def test_batched_stripes_equal_single_stripes(ctx):
    """clrt_render_stripes_batch: 1..4 stripes (one camera each) by one pair of launches == the same stripes rendered one by one."""
    import torch
    from openclraytracer_b200 import frame
    buf = scenes.load_bundled("tenryuu")
    scene, bvh = build_gpu(ctx, buf)
    W, H = 321, 203
    cams = [frame.orbit_camera(buf, W, H, s) for s in range(4)]
    for nranks, rank in ((1, 0), (3, 1), (8, 7)):
        cap, tiles = api.frame_stripe_pixels(W, H, rank, nranks)
        units = tiles * 256
        single = api.Buffer(ctx, max(cap, 1) * 4)
        want = []
        for c in cams:
            bvh.render_stripe(c, True, rank, nranks, single.getCLMem())
            want.append(single.copyToHost(np.uint32, cap)[:units].copy())
        for k in (1, 2, 4):
            batch = api.Buffer(ctx, max(units, 1) * 4 * k)
            bvh.render_stripes_batch(ctx.stream, 0, cams[:k], rank, nranks, batch.getCLMem())
            ctx.finish()
            got = batch.copyToHost(np.uint32, units * k)
            for f in range(k):
                assert np.array_equal(got[f * units:(f + 1) * units], want[f]), (nranks, rank, k, f)
            batch.free()
        single.free()
    bvh.check()
    bvh.free(); scene.free()


def test_full_size_soup_1m_properties(ctx):
    """BASELINE config 3 size: 1M-triangle soup.  Size-independent properties: sorted keys, leaf permutation,
    hierarchy + containment, FAST == REFERENCE_ORDER closest hits, idempotent rebuild."""
    n = 1_000_000
    buf = scenes.soup_scene(n)
    scene, bvh = build_gpu(ctx, buf)
    pairs, nodes = bvh.export_pairs(), bvh.export_nodes()
    assert (pairs["x"][1:] >= pairs["x"][:-1]).all()
    eq = pairs["x"][1:] == pairs["x"][:-1]
    assert (pairs["y"][1:][eq] > pairs["y"][:-1][eq]).all()
    assert np.array_equal(np.sort(pairs["y"]), np.arange(n, dtype=np.uint32))
    check_hierarchy(nodes, n)
    bvh.initializeFrame(); bvh.construct()
    assert digest(masked_nodes(bvh.export_nodes(), n)) == digest(masked_nodes(nodes, n))
    cam = api.demo_camera(buf, 640, 360)
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER); bvh.generateContacts(cam); a = bvh.download_primary_contacts()
    bvh.set_trace_mode(api.TRACE_FAST); bvh.generateContacts(cam); b = bvh.download_primary_contacts()
    assert (a["nd"][:, 3] > 0).mean() > 0.2
    compare_fast(b, a, None)
    bvh.free(); scene.free()


def test_errors_are_reported_not_swallowed(ctx):
    scene = api.Scene(np.zeros(200, dtype=np.uint8))
    with pytest.raises(api.ClrtError, match="totalDataSize"):
        scene.loadToGPU(ctx)
    buf = scenes.cube_scene()
    s = api.Scene(buf); s.loadToGPU(ctx)
    bvh = api.BVHManager(ctx, s); bvh.initialize()
    with pytest.raises(api.ClrtError, match="initializeFrame"):
        bvh.construct()
    bvh.initializeFrame()
    with pytest.raises(api.ClrtError, match="not constructed|was not constructed"):
        bvh.generateContacts(api.demo_camera(buf, 8, 8))
    bvh.free(); s.free()


@pytest.mark.parametrize("name", ["tenryuu", "clumsydragon", "soup"])
def test_fused_shadow_pass_equals_three_kernel_pipeline(ctx, name):
    """The frame's fused shadow+shade kernel (occlusion queries) gives the pixels and ray counts of
    generate_shadow_rays -> trace_rays -> shade, bit for bit."""
    buf = scenes.soup_scene(60_000, seed=5) if name == "soup" else scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    for cam in (api.demo_camera(buf, 320, 180), api.demo_camera(buf, 97, 61)):
        bvh.set_fused_shadows(False)
        b, rb = bvh.render_frame_host(cam, shadows=True, want_rays=True)
        b = b.copy()
        for mode in (1, 2):                     # 1: primary kernel + fused shadow/shade kernel (default); 2: the one-kernel frame
            bvh.set_fused_shadows(mode)
            a, ra = bvh.render_frame_host(cam, shadows=True, want_rays=True)
            assert np.array_equal(a, b) and ra == rb, mode
        assert (b[..., 3] == 255).any()
    bvh.set_fused_shadows(True)
    # stand-alone entry point over the primary contacts
    cam = api.demo_camera(buf, 160, 90)
    n = 160 * 90
    bvh.generateContacts(cam)
    u8a = api.Buffer(ctx, n * 4); u8b = api.Buffer(ctx, n * 4)
    bvh.shadow_shade(cam, bvh.getPrimaryContacts(), n, u8a.getCLMem())
    L1 = scene.num_lights() + 1
    srays = api.Buffer(ctx, n * L1 * 48); scont = api.Buffer(ctx, n * L1 * 32)
    api.generate_shadow_rays(ctx, scene, cam, bvh.getPrimaryContacts(), n, srays.getCLMem())
    bvh.generateContacts(srays, scont, n * L1)
    api.shade(ctx, scene, cam, bvh.getPrimaryContacts(), n, scont.getCLMem(), u8b.getCLMem(), None)
    assert np.array_equal(u8a.copyToHost(np.uint32), u8b.copyToHost(np.uint32))
    for x in (u8a, u8b, srays, scont):
        x.free()
    bvh.free(); scene.free()


@pytest.mark.parametrize("name", ["cube", "gardenlight", "clumsydragon", "soup"])
def test_trace_tree_returns_the_reference_tree_hits(ctx, O, name):
    """The separate trace tree (trace_tree.cu) is another hierarchy over the same triangles: FAST contacts through it equal
    FAST contacts through the reference tree — ids included (ties go to the reference's winner) — and both meet the oracle."""
    buf = scenes.soup_scene(30_000, seed=7, tris_per_mesh=4096) if name == "soup" else scenes.load_bundled(name)
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    bvh = api.BVHManager(ctx, scene); bvh.initialize()
    b = O.bvh_build(buf)
    cam = api.demo_camera(buf, 400, 300)
    want = O.bvh_trace_primary(cam, b)
    got = {}
    for policy in (api.TRACE_TREE_NEVER, api.TRACE_TREE_ALWAYS, api.TRACE_TREE_AUTO):
        bvh.set_trace_tree(policy)
        bvh.initializeFrame(); bvh.construct()
        if policy != api.TRACE_TREE_AUTO:
            assert bvh.has_trace_tree() == (policy == api.TRACE_TREE_ALWAYS and scene.triangle_count() >= 2)
        assert np.array_equal(bvh.export_pairs(), b["pairs"])            # the reference-layout build is untouched
        bvh.set_trace_mode(api.TRACE_FAST)
        bvh.generateContacts(cam)
        got[policy] = bvh.download_primary_contacts()
        compare_fast(got[policy], want, b["tris"])
    a, c = got[api.TRACE_TREE_NEVER], got[api.TRACE_TREE_ALWAYS]
    assert np.array_equal(a["pad"][:, 0], c["pad"][:, 0]) and np.array_equal(a["nd"], c["nd"])
    if name == "clumsydragon":
        assert bvh.has_trace_tree()          # auto: long equal-key runs + a scene box below zero
    bvh.free(); scene.free()


def test_frame_renderer_pipeline_equals_synchronous_frames(ctx):
    """FrameRenderer keeps two frames in flight (second stream for untile + read-back); every frame must equal the one the
    synchronous C-ABI frame call returns for the same camera."""
    from openclraytracer_b200 import frame
    buf = scenes.load_bundled("tenryuu")
    r = frame.FrameRenderer(buf, 320, 200, "BVH", 0, 1, 0)
    r.build()
    cams = [frame.orbit_camera(buf, 320, 200, s) for s in range(5)]
    want = [r.accel.render_frame_host(c, True)[0].copy() for c in cams]
    outs = []
    for c in cams[:2]:
        outs.append(r.render_host_async(c, True))
    r.finish()
    got01 = [o.numpy().reshape(200, 320).copy() for o in outs]
    for k in range(2):
        assert np.array_equal(got01[k].view(np.uint8).reshape(200, 320, 4), want[k])
    for k in range(2, 5):
        got = r.render_host(cams[k], True).numpy().reshape(200, 320).view(np.uint8).reshape(200, 320, 4)
        assert np.array_equal(got, want[k])
        r.render_device(cams[k], True); r.finish()
        assert np.array_equal(r.raster.cpu().numpy().view(np.uint8).reshape(200, 320, 4), want[k])
    r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("exchange", ["push", "fused"])
def test_frame_renderer_peer_memory_exchange_equals_synchronous_frames(ctx, exchange, monkeypatch):
    """The exchanges over peer memory (csrc/peer.cu; here with one GPU, i.e. the GPU is its own only peer): pixels stored in
    raster order by the push pass / by the shading kernel, ready + ack words per frame slot.  More frames than slots, so that
    every slot is reused; each frame must equal the synchronous C-ABI frame.  One frame is 321x203 (ragged tiles)."""
    from openclraytracer_b200 import frame
    monkeypatch.setenv("CLRT_EXCHANGE", exchange)
    buf = scenes.load_bundled("tenryuu")
    for res in ((320, 200), (321, 203)):
        r = frame.FrameRenderer(buf, res[0], res[1], "BVH", 0, 1, 0)
        assert r.exchange == exchange
        r.build()
        cams = [frame.orbit_camera(buf, res[0], res[1], s) for s in range(2 * r.DEPTH + 1)]
        want = [r.accel.render_frame_host(c, True)[0].copy() for c in cams]
        for k, c in enumerate(cams):
            r.render_device(c, True); r.finish()
            assert np.array_equal(r.raster.cpu().numpy().view(np.uint8).reshape(res[1], res[0], 4), want[k]), (res, k)
        outs = [r.render_host_async(c, True) for c in cams[:r.DEPTH]]
        r.finish()
        for k, o in enumerate(outs):
            assert np.array_equal(o.numpy().view(np.uint8).reshape(res[1], res[0], 4), want[k]), (res, k)
        got = r.render_host(cams[-1], False).numpy().view(np.uint8).reshape(res[1], res[0], 4)      # no shadows: the push pass in both modes
        assert np.array_equal(got, r.accel.render_frame_host(cams[-1], False)[0])
        r.close()


@pytest.mark.gpu
@pytest.mark.parametrize("ss", [2, 3])
def test_supersampled_frame_is_the_box_filtered_fine_frame(ctx, ss):
    """Camera.supersampilngFactor (CLStructs.h:49, unused by the reference; its TODO 2.4): s*s rays per pixel = the frame of the
    camera (s*resX, s*resY, s*FOVDistance) averaged over s x s blocks with rounding; bit-exact against numpy on the fine frame."""
    buf = scenes.load_bundled("tenryuu")
    scene, bvh = build_gpu(ctx, buf)
    W, H = 161, 100
    cam = api.demo_camera(buf, W, H)
    fine = api.make_camera(W * ss, H * ss, np.float32(cam["FOVDistance"]) * np.float32(ss), cam["viewTransform"])
    cam_ss = cam.copy(); cam_ss["supersampilngFactor"] = ss
    for shadows in (False, True):
        got, rays = bvh.render_frame_host(cam_ss, shadows, want_rays=True)
        hi, rays_hi = bvh.render_frame_host(fine, shadows, want_rays=True)
        want = ((hi.reshape(H, ss, W, ss, 4).astype(np.uint32).sum(axis=(1, 3)) + ss * ss // 2) // (ss * ss)).astype(np.uint8)
        assert got.shape == (H, W, 4) and np.array_equal(got, want)
        assert rays == rays_hi and rays >= W * H * ss * ss
        assert 0 < (got[..., 3] % 255 != 0).sum() < W * H // 4          # partial coverage exists, and only along silhouettes
    cam0 = cam.copy(); cam0["supersampilngFactor"] = 0
    assert np.array_equal(bvh.render_frame_host(cam0, True)[0], bvh.render_frame_host(cam, True)[0])
    cam9 = cam.copy(); cam9["supersampilngFactor"] = 9
    with pytest.raises(api.ClrtError):
        bvh.render_frame_host(cam9, True)
    bvh.free(); scene.free()


@pytest.mark.gpu
@pytest.mark.parametrize("kind", __import__("edge_scenes").KINDS)
def test_edge_scenes_build_and_trace_vs_oracle(ctx, O, kind):
    """tests/edge_scenes.py (zero extent on an axis, all-negative box, identical triangles, zero-area triangles, collapsed axes,
    clustered keys): triangles, Morton pairs and nodes bit-exact, REFERENCE_ORDER contacts bit-exact, FAST within the north-star
    tolerances — through whichever trace structure construct() chose, and through both when forced."""
    import edge_scenes
    buf = edge_scenes.edge_scene(kind)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    n = len(b["tris"])
    assert np.array_equal(scene.export_triangles(), b["tris"])
    assert np.array_equal(bvh.export_pairs(), b["pairs"])
    nodes = bvh.export_nodes()
    assert np.array_equal(masked_nodes(nodes, n), masked_nodes(b["nodes"], n)[: 2 * n - 1])
    cam = edge_scenes.edge_camera(kind, 96, 64)
    want = O.bvh_trace_primary(cam, b)
    hit = want["nd"][:, 3] > 0
    assert hit.sum() > 10
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    got = bvh.download_primary_contacts()
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    assert np.array_equal(got["nd"][hit], want["nd"][hit]) and np.array_equal(got["pad"][hit, 0], want["pad"][hit, 0])
    bvh.set_trace_mode(api.TRACE_FAST)
    for policy in (api.TRACE_TREE_AUTO, api.TRACE_TREE_ALWAYS, api.TRACE_TREE_NEVER):
        bvh.set_trace_tree(policy)
        bvh.initializeFrame(); bvh.construct()
        bvh.generateContacts(cam)
        compare_fast(bvh.download_primary_contacts(), want, b["tris"])
    bvh.free(); scene.free()
