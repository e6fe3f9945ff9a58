"""GPU: the C++ host mirror (include/clrt/CLRayTracer.hpp, host/) driven the way a caller of the reference drives it —
CLExecutionContext / CLBuffer / BitonicSort / PrefixSum / Scene / BVHManager / TwoLevelGridManager with Result + Errata —
compared with the oracle; and the headless demo entry point with the reference's command line."""
import os
import subprocess
import numpy as np
import pytest
from conftest import ROOT
from openclraytracer_b200 import api, scenes, layouts as L

pytestmark = pytest.mark.gpu
BUILD = os.path.join(ROOT, "host", "build")
OBJ_DIR = os.path.join(ROOT, "tests", "golden", "obj")


@pytest.fixture(scope="module")
def host_binaries():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "host")], check=True)
    return BUILD


def run(cmd, **kw):
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=300, **kw)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    return r.stdout


@pytest.mark.parametrize("name", ["cube", "tenryuu"])
def test_class_interface_contacts_match_the_oracle(host_binaries, O, tmp_path, name):
    buf = scenes.load_bundled(name)
    (tmp_path / "scene.bin").write_bytes(buf.tobytes())
    out = run([os.path.join(host_binaries, "host_selftest"), str(tmp_path / "scene.bin"), "160", "120", str(tmp_path / "o")])
    assert "host_selftest ok" in out
    assert np.array_equal(np.fromfile(tmp_path / "o_scene.bin", dtype=np.uint8), buf)
    cam = np.fromfile(tmp_path / "o_camera.bin", dtype=L.CAMERA)[0]
    want_cam = api.demo_camera(buf, 160, 120)
    assert cam["FOVDistance"] == want_cam["FOVDistance"] and np.array_equal(cam["viewTransform"], want_cam["viewTransform"])
    b = O.bvh_build(buf)
    want = O.bvh_trace_primary(cam, b)
    got = np.fromfile(tmp_path / "o_bvh_contacts.bin", dtype=L.CONTACT)
    hit = want["nd"][:, 3] > 0
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    assert np.array_equal(got["nd"][hit], want["nd"][hit]) and np.array_equal(got["pad"][hit, 0], want["pad"][hit, 0])
    if O.have_ref():                                   # grid contacts against the reference's own TwoLevelGrid.h
        g = O.RefTlg(buf)
        wg = g.trace_primary(cam)
        gg = np.fromfile(tmp_path / "o_tlg_contacts.bin", dtype=L.CONTACT)
        assert np.array_equal(gg["nd"], wg["nd"])
        g.free()


def test_scene_load_through_the_class_interface(host_binaries, tmp_path):
    """Scene::load of a scene file with an OBJ fixture through the classes: the packed bytes equal the committed
    reference-loader bytes and the BVH path runs on them."""
    with open(tmp_path / "s.txt", "w") as f:
        f.write("LIGHT 0.0 10.0 -330.0 1000.0\nMESH " + os.path.join(OBJ_DIR, "quirks.obj"))
    out = run([os.path.join(host_binaries, "host_selftest"), str(tmp_path / "s.txt"), "96", "64", str(tmp_path / "o"), "bvh-only"])     # 13 scattered triangles: the grid's resolution formula degenerates (reported as Error)
    assert "host_selftest ok" in out
    want = np.load(os.path.join(OBJ_DIR, "packed.npz"))["quirks"]
    assert np.array_equal(np.fromfile(tmp_path / "o_scene.bin", dtype=np.uint8), want)


def test_headless_demo_command_line(host_binaries, ctx, tmp_path):
    """The reference's switches (-winH -winW -accStruct -scene, -headersPath ignored) + a camera script; frame 0 equals the
    frame the Python mirror renders for the demo camera through the same structure (same library, other host language)."""
    with open(tmp_path / "s.txt", "w") as f:
        f.write("LIGHT 1.5 2.5 -3.5 40.0\nLIGHT 0.0 0.0 0.0 1.0\nMESH " + os.path.join(OBJ_DIR, "nomtl.obj"))
    buf = np.load(os.path.join(OBJ_DIR, "packed.npz"))["nomtl"]
    cam = api.demo_camera(buf, 120, 90)
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    for acc in ("BVH", "GRID"):
        out = run([os.path.join(host_binaries, "OpenCLRayTracerDemo"), "-winH", "90", "-winW", "120", "-accStruct", acc, "-headersPath", "x",
                   "-scene", str(tmp_path / "s.txt"), "-script", "UaLw", "-out", str(tmp_path / acc)])
        assert out.count("pixels hit") == 5
        raw = (tmp_path / f"{acc}_0000.ppm").read_bytes()
        assert raw.startswith(b"P6\n120 90\n255\n")
        img = np.frombuffer(raw[len(b"P6\n120 90\n255\n"):], dtype=np.uint8).reshape(90, 120, 3)
        assert img.any()
        m = api.BVHManager(ctx, scene) if acc == "BVH" else api.TwoLevelGridManager(ctx, scene)
        m.initialize(); m.initializeFrame(); m.construct()
        m.generateContacts(cam)
        rgba = api.Buffer(ctx, 120 * 90 * 4)
        api.shade(ctx, scene, cam, m.getPrimaryContacts(), 120 * 90, None, rgba.getCLMem())
        want = rgba.copyToHost(np.uint8, 120 * 90 * 4).reshape(90, 120, 4)[:, :, :3]
        assert np.array_equal(img, want)
        moved = np.frombuffer((tmp_path / f"{acc}_0004.ppm").read_bytes()[len(b"P6\n120 90\n255\n"):], dtype=np.uint8).reshape(90, 120, 3)
        assert not np.array_equal(moved, img)              # the script moved the camera
        rgba.free(); m.free()
    scene.free()
    assert "Command line parameters" in run([os.path.join(host_binaries, "OpenCLRayTracerDemo"), "-winH", "10"])
