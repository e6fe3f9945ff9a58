"""GPU: the KD-tree manager (csrc/kd.cu; row f4, the reference's TODO 2.1) through the same C ABI handle and the same harness as the
other two structures.  The reference has no KD-tree to be bit-exact with; what a KD-tree must return is the closest hit, so its
contacts are compared with the ORACLE's BVH contacts by the north-star tolerances (ids >= 99.99 %, t 1e-4, normals exact where the
ids agree) — camera rays, generic rays, incoherent bounce rays, the edge scenes, a frame through the three-call shadow pipeline."""
import os
import subprocess
import numpy as np
import pytest
from conftest import ROOT
from openclraytracer_b200 import api, scenes, frame, layouts as L
from test_gpu_bvh import compare_fast

pytestmark = pytest.mark.gpu


def build_kd(ctx, buf):
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    kd = api.KDTreeManager(ctx, scene); kd.initialize(); kd.initializeFrame(); kd.construct()
    return scene, kd


@pytest.mark.parametrize("name", ["cube", "gardenlight", "tenryuu", "manfigure", "clumsydragon", "soup"])
def test_kd_contacts_are_the_bvh_closest_hits(ctx, O, name):
    buf = scenes.soup_scene(60_000, seed=13, tris_per_mesh=8192) if name == "soup" else scenes.load_bundled(name)
    scene, kd = build_kd(ctx, buf)
    info = kd.info()
    n = scene.triangle_count()
    assert info["leaves"] >= 1 and info["pairs"] >= n and 1 <= info["levels"] <= 10
    b = O.bvh_build(buf)
    for cam in (api.demo_camera(buf, 320, 200), api.demo_camera(buf, 97, 61), frame.orbit_camera(buf, 256, 144, 9)):
        want = O.bvh_trace_primary(cam, b)
        kd.generateContacts(cam)
        got = kd.download_primary_contacts()
        compare_fast(got, want, b["tris"])
        hit = want["nd"][:, 3] > 0
        assert np.array_equal(got["pixelIndex"], want["pixelIndex"]) and np.array_equal(got["materialIndex"][hit], want["materialIndex"][hit])
    kd.free(); scene.free()


def test_kd_generic_and_bounce_rays(ctx, O):
    buf = scenes.soup_scene(30_000, seed=17, tris_per_mesh=4096)
    scene, kd = build_kd(ctx, buf)
    b = O.bvh_build(buf)
    W, H = 192, 108
    n = W * H
    cam = api.demo_camera(buf, W, H)
    kd.generateContacts(cam)
    rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_bounce_rays(ctx, cam, kd.getPrimaryContacts(), n, 20161116, rays.getCLMem())
    r = rays.copyToHost(L.RAY, n)
    r["direction"][7] = 0                                          # a zero direction is a miss
    rays.upload(r)
    kd.generateContacts(rays, contacts, n)
    got = contacts.copyToHost(L.CONTACT, n)
    live = (r["direction"][:, :3] != 0).any(axis=1)
    want = np.zeros(n, dtype=L.CONTACT)
    want[live] = O.bvh_trace_rays(np.ascontiguousarray(r[live]), b)
    assert (got["nd"][~live, 3] == 0).all() and (want["nd"][live, 3] > 0).sum() > 100
    g2, w2 = got[live].copy(), want[live]
    compare_fast(g2, w2, b["tris"])
    rays.free(); contacts.free(); kd.free(); scene.free()


@pytest.mark.parametrize("kind", __import__("edge_scenes").KINDS)
def test_kd_edge_scenes(ctx, O, kind):
    """Zero extent on an axis, all-negative boxes, identical triangles, zero-area triangles, collapsed axes, clustered keys."""
    import edge_scenes
    buf = edge_scenes.edge_scene(kind)
    scene, kd = build_kd(ctx, buf)
    b = O.bvh_build(buf)
    cam = edge_scenes.edge_camera(kind, 96, 64)
    want = O.bvh_trace_primary(cam, b)
    kd.generateContacts(cam)
    compare_fast(kd.download_primary_contacts(), want, b["tris"])
    kd.free(); scene.free()


def test_kd_rebuild_is_idempotent_and_errors_are_reported(ctx):
    buf = scenes.load_bundled("tenryuu")
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    kd = api.KDTreeManager(ctx, scene); kd.initialize()
    with pytest.raises(api.ClrtError, match="initializeFrame"):
        kd.construct()
    kd.initializeFrame()
    with pytest.raises(api.ClrtError, match="not constructed"):
        kd.generateContacts(api.demo_camera(buf, 8, 8))
    kd.construct()
    cam = api.demo_camera(buf, 160, 100)
    kd.generateContacts(cam); a = kd.download_primary_contacts(); i1 = kd.info()
    kd.initializeFrame(); kd.construct()
    kd.generateContacts(cam); b = kd.download_primary_contacts()
    assert kd.info() == i1 and a.tobytes() == b.tobytes()
    kd.free(); scene.free()


def test_kd_frame_with_shadows_equals_the_bvh_frame(ctx):
    """render_frame_host on a KD-tree (three-call shadow pipeline through the generic-ray entry) gives the BVH's frame, up to the
    pixels where an exact tie picked another triangle."""
    buf = scenes.load_bundled("tenryuu")
    scene, kd = build_kd(ctx, buf)
    bvh = api.BVHManager(ctx, scene); bvh.initialize(); bvh.initializeFrame(); bvh.construct()
    cam = api.demo_camera(buf, 320, 200)
    a, ra = kd.render_frame_host(cam, shadows=True, want_rays=True)
    b, rb = bvh.render_frame_host(cam, shadows=True, want_rays=True)
    diff = (np.abs(a.astype(np.int32) - b.astype(np.int32)).max(axis=2) > 1).sum()
    print(f"[kd frame] pixels beyond 1 LSB of the BVH frame: {int(diff)} of {a.shape[0] * a.shape[1]}")
    assert diff <= 3 and ra == rb
    kd.free(); bvh.free(); scene.free()


def test_demo_accepts_the_kd_tree(tmp_path):
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "host")], check=True)
    buf = scenes.load_bundled("gardenlight")
    (tmp_path / "scene.bin").write_bytes(buf.tobytes())
    outs = {}
    for acc in ("BVH", "KD"):
        r = subprocess.run([os.path.join(ROOT, "host", "build", "OpenCLRayTracerDemo"), "-winH", "120", "-winW", "160", "-accStruct", acc,
                            "-scene", str(tmp_path / "scene.bin"), "-out", str(tmp_path / acc)], capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        outs[acc] = np.frombuffer((tmp_path / f"{acc}_0000.ppm").read_bytes()[len(b"P6\n160 120\n255\n"):], dtype=np.uint8).astype(np.int32)
    assert outs["BVH"].any() and (np.abs(outs["BVH"] - outs["KD"]) > 1).sum() <= 6
