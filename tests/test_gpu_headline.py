"""GPU parity on BASELINE.json's OWN configurations, CUDA path vs the ORACLE (not vs itself):

  configs[2]  Soup-1M (seed 20161116): Morton pairs + nodes bit-exact; REFERENCE_ORDER contacts bit-exact and FAST contacts within
              the north-star tolerances at 960x540 on two poses of the bench camera path; the frame's fused shadow_shade_kernel
              RGBA8 vs oracle shadow rays -> oracle traversal -> oracle Blinn-Phong within 1 LSB.
  configs[3]  Spheres-10M: Morton pairs + nodes bit-exact (digests) vs the oracle's flattened accessor; >= 100k-ray samples of the
              4K primary pass and of the diffuse-bounce pass vs the oracle on the SAME rays.
  frame path  the FrameRenderer's frames (frame slots; peer-memory exchange with the GPU as its own peer) vs the oracle frame.
  configs[1]  the shipped Bin/*.txt scenes whose models exist (ClumsyDragon, man, tenryuu; skull / swfighter are missing blobs)
              through the headless Demo at the .bat files' 512x512, frame 0 vs the oracle's host loop
              (R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:298-316).

Every test prints the count of mismatching hit ids per configuration (north star: >= 99.99 % agreement, mismatches only on
near ties |dt| <= 1e-5 t)."""
import os
import subprocess
import numpy as np
import pytest
from conftest import ROOT, masked_nodes, digest
from openclraytracer_b200 import api, scenes, frame, layouts as L
from test_gpu_bvh import build_gpu, compare_fast, check_hierarchy

pytestmark = pytest.mark.gpu
SEED = 20161116


def id_mismatches(got, want):
    hit = want["nd"][:, 3] > 0
    return int((got["pad"][hit, 0] != want["pad"][hit, 0]).sum()), int(hit.sum())


def oracle_frame(O, cam, buf, b, contacts=None):
    """The oracle's frame: primary contacts -> shadow rays (live ones only are traced) -> shadow contacts -> Blinn-Phong RGBA8."""
    c = O.bvh_trace_primary(cam, b) if contacts is None else contacts
    sr = O.shadow_rays(cam, buf, c)
    live = sr["idx"] != 0xFFFFFFFF
    sc = np.zeros(len(sr), dtype=L.CONTACT)
    sc[live] = O.bvh_trace_rays(np.ascontiguousarray(sr[live]), b)
    return c, sr, sc, O.shade(cam, buf, c, shadow=sc)


def assert_rgba_within_1lsb(got_u32, want_u32, what):
    """<= 1 LSB everywhere, except pixels whose shadow test is a near tie (the occluder's distance within 1e-5 of the distance to
    the light, where an occlusion query and a closest-hit query may round apart); those are counted and must be < 1e-5 of the frame."""
    g = got_u32.view(np.uint8).reshape(-1, 4).astype(np.int32)
    w = want_u32.view(np.uint8).reshape(-1, 4).astype(np.int32)
    bad = np.abs(g - w).max(axis=1) > 1
    print(f"[{what}] pixels beyond 1 LSB: {int(bad.sum())} of {len(bad)}")
    assert bad.sum() <= max(1, int(1e-5 * len(bad))), what


@pytest.fixture(scope="module")
def soup1m(ctx, O):
    buf = scenes.soup_scene(1_000_000)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    yield buf, scene, bvh, b
    bvh.free(); scene.free()


def test_soup1m_build_bit_exact_vs_oracle(soup1m):
    buf, scene, bvh, b = soup1m
    n = 1_000_000
    assert np.array_equal(scene.export_triangles(), b["tris"])
    assert np.array_equal(bvh.export_pairs(), b["pairs"])                                   # Morton codes + sorted order
    nodes = bvh.export_nodes()
    assert digest(masked_nodes(nodes, n)) == digest(masked_nodes(b["nodes"], n)[: 2 * n - 1])   # topology + boxes
    check_hierarchy(nodes, n)


@pytest.mark.parametrize("pose", [3, 11])
def test_soup1m_contacts_vs_oracle_960x540(soup1m, O, pose):
    buf, scene, bvh, b = soup1m
    cam = frame.orbit_camera(buf, 960, 540, pose)
    want = O.bvh_trace_primary(cam, b)
    hit = want["nd"][:, 3] > 0
    assert 0.2 < hit.mean() < 0.9
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    got = bvh.download_primary_contacts()
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    assert np.array_equal(got["nd"][hit], want["nd"][hit]) and np.array_equal(got["pad"][hit, 0], want["pad"][hit, 0])
    bvh.set_trace_mode(api.TRACE_FAST)
    bvh.generateContacts(cam)
    fast = bvh.download_primary_contacts()
    compare_fast(fast, want, b["tris"])
    print("[soup1m pose %d] FAST id mismatches: %d of %d hits" % ((pose,) + id_mismatches(fast, want)))


def test_soup1m_fused_shadow_shade_vs_oracle(soup1m, ctx, O):
    """shadow_shade_kernel (what the bench's frame runs) against oracle shadow rays + oracle traversal + oracle shading."""
    buf, scene, bvh, b = soup1m
    W, H = 960, 540
    cam = frame.orbit_camera(buf, W, H, 5)
    c, sr, sc, want = oracle_frame(O, cam, buf, b)
    assert (sc["nd"][:, 3] > 0).mean() > 0.05                        # lights are actually occluded somewhere
    bvh.set_trace_mode(api.TRACE_FAST)
    bvh.generateContacts(cam)
    u8 = api.Buffer(ctx, W * H * 4)
    bvh.shadow_shade(cam, bvh.getPrimaryContacts(), W * H, u8.getCLMem())
    assert_rgba_within_1lsb(u8.copyToHost(np.uint32), want, "soup1m fused shadow+shade")
    img, rays = bvh.render_frame_host(cam, shadows=True, want_rays=True)           # the frame entry point: same pixels
    assert_rgba_within_1lsb(np.ascontiguousarray(img).view(np.uint32).reshape(-1), want, "soup1m render_frame_host")
    assert rays == W * H + int((sr["idx"] != 0xFFFFFFFF).sum())
    u8.free()


@pytest.mark.parametrize("exchange", ["nccl", "fused", "fused-batch4", "fused-batch2"])
def test_frame_renderer_frames_vs_oracle(ctx, O, exchange, monkeypatch):
    """The bench's frame path (FrameRenderer: frame slots on several streams; with CLRT_EXCHANGE=fused the shading kernel stores
    the pixels through the peer-memory block, the GPU being its own peer) against the ORACLE's frame, more frames than slots."""
    if "-batch" in exchange:                     # several frames per pair of launches (clrt_render_frames_peers), as at 4 / 8 GPUs
        monkeypatch.setenv("CLRT_FRAME_BATCH", exchange[-1])
        exchange = "fused"
    monkeypatch.setenv("CLRT_EXCHANGE", exchange)
    buf = scenes.soup_scene(150_000, seed=9)
    b = O.bvh_build(buf)
    W, H = 480, 270
    r = frame.FrameRenderer(buf, W, H, "BVH", 0, 1, 0)
    assert r.exchange == exchange
    r.build()
    poses = list(range(r.DEPTH + 3))             # more frames than slots, and a last batch that does not fill up
    outs = []
    for p in poses:
        outs.append(r.render_host_async(frame.orbit_camera(buf, W, H, p), True))
        if len(outs) == r.DEPTH:                                   # slots are about to be reused: collect what is in flight
            r.finish()
            outs = [o.numpy().copy() if hasattr(o, "numpy") else o for o in outs]
    r.finish()
    outs = [o.numpy().copy() if hasattr(o, "numpy") else o for o in outs]
    for p, got in zip(poses, outs):
        cam = frame.orbit_camera(buf, W, H, p)
        c, sr, sc, want = oracle_frame(O, cam, buf, b)
        assert_rgba_within_1lsb(np.ascontiguousarray(got).view(np.uint32).reshape(-1), want, f"frame renderer {exchange} pose {p}")
    r.close()


# ---- configs[3]: Spheres-10M ---------------------------------------------------------------------------------------------------
@pytest.fixture(scope="module")
def spheres10m(ctx, O):
    buf = scenes.sphere_field_scene()
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)                        # flattened accessor (oracle_flatten), validated against the literal walk on small scenes
    yield buf, scene, bvh, b
    bvh.free(); scene.free()


def test_spheres10m_build_bit_exact_vs_oracle(spheres10m):
    buf, scene, bvh, b = spheres10m
    n = 10_000_000
    assert scene.triangle_count() == n
    assert digest(scene.export_triangles()) == digest(b["tris"])
    assert digest(bvh.export_pairs()) == digest(b["pairs"])
    assert digest(masked_nodes(bvh.export_nodes(), n)) == digest(masked_nodes(b["nodes"], n)[: 2 * n - 1])


def test_spheres10m_primary_and_bounce_samples_vs_oracle(spheres10m, ctx, O):
    """4K primary rays + one diffuse bounce per hit (configs[3]); a strided sample of >= 100k rays of each pass is traced by the
    oracle on the same rays (generic-ray entry) and compared by the north-star tolerances."""
    buf, scene, bvh, b = spheres10m
    W, H = 3840, 2160
    n = W * H
    cam = api.demo_camera(buf, W, H)
    bvh.set_trace_mode(api.TRACE_FAST)
    bvh.generateContacts(cam)
    prim = bvh.download_primary_contacts()
    prays = api.Buffer(ctx, n * 48); rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_primary_rays(ctx, cam, prays.getCLMem())
    api.generate_bounce_rays(ctx, cam, bvh.getPrimaryContacts(), n, SEED, rays.getCLMem())
    bvh.generateContacts(rays, contacts, n)
    bounce = contacts.copyToHost(L.CONTACT, n)
    sub = np.arange(0, n, 61, dtype=np.int64)                                      # 135 987 rays
    pr = np.ascontiguousarray(prays.copyToHost(L.RAY, n)[sub])
    want = O.bvh_trace_rays(pr, b)
    got = prim[sub].copy(); got["pixelIndex"] = want["pixelIndex"]
    compare_fast(got, want, b["tris"])
    print("[spheres10m primary] FAST id mismatches: %d of %d hits" % id_mismatches(got, want))
    br_all = rays.copyToHost(L.RAY, n)
    live = np.flatnonzero(br_all["idx"] != 0xFFFFFFFF)
    sub2 = live[:: max(1, len(live) // 120_000)]
    br = np.ascontiguousarray(br_all[sub2])
    want2 = O.bvh_trace_rays(br, b)
    compare_fast(bounce[sub2], want2, b["tris"])
    print("[spheres10m bounce] FAST id mismatches: %d of %d hits (%d rays)" % (id_mismatches(bounce[sub2], want2) + (len(sub2),)))
    assert len(sub) >= 100_000 and len(sub2) >= 100_000
    for x in (prays, rays, contacts):
        x.free()


# ---- configs[1] / row f3: the shipped demo scenes through the headless Demo vs the oracle's host loop ---------------------------
@pytest.mark.parametrize("name", ["clumsydragon", "manfigure", "tenryuu"])      # Bin/ClumsyDragon.txt, man.txt, tenryuu.txt
def test_demo_frame0_of_shipped_scenes_vs_oracle_host_loop(O, tmp_path, name):
    """Bin/<scene>.bat runs `OpenCLRayTracerDemo -winH 512 -winW 512 -accStruct BVH -scene <scene>.txt`; the scene's packed bytes
    are the committed fixture (tests/golden/scenes, packed by the reference's loader from that very .txt: LIGHT 0 10 -330 1000 + the
    model; tests/test_host_scene.py pins Scene::load of the .txt to the same bytes).  Frame 0 (no shadows, like the reference)
    must equal the oracle's host loop: bvh_generate_contact per pixel -> cpuShadeBlinnPhong -> RGBA8, within 1 LSB."""
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "host")], check=True)
    buf = scenes.load_bundled(name)
    (tmp_path / "scene.bin").write_bytes(buf.tobytes())
    W = H = 512
    r = subprocess.run([os.path.join(ROOT, "host", "build", "OpenCLRayTracerDemo"), "-winH", str(H), "-winW", str(W), "-accStruct", "BVH",
                        "-headersPath", "..\\..\\include", "-scene", str(tmp_path / "scene.bin"), "-out", str(tmp_path / "f")],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    raw = (tmp_path / "f_0000.ppm").read_bytes()
    head = b"P6\n%d %d\n255\n" % (W, H)
    assert raw.startswith(head)
    img = np.frombuffer(raw[len(head):], dtype=np.uint8).reshape(H * W, 3).astype(np.int32)
    cam = O.demo_camera(buf, W, H)
    b = O.bvh_build(buf)
    c = O.bvh_trace_primary(cam, b)
    if O.have_ref():
        col = O.ref_shade(cam, c, buf)                                      # the reference's own cpuShadeBlinnPhong
        want = np.floor(np.clip(col, 0, 1) * 255.0 + 0.5).astype(np.int32)[:, :3]
        want[c["nd"][:, 3] <= 0] = 0
    else:
        want = O.shade(cam, buf, c).view(np.uint8).reshape(-1, 4)[:, :3].astype(np.int32)
    bad = np.abs(img - want).max(axis=1) > 1
    print(f"[demo {name}] pixels beyond 1 LSB: {int(bad.sum())} of {len(bad)}; hit pixels {int((c['nd'][:, 3] > 0).sum())}")
    assert (c["nd"][:, 3] > 0).sum() > 1000
    assert bad.sum() <= 2                      # FAST traversal: an exact-tie pixel may take the other triangle's normal
