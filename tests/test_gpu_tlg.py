"""GPU parity: Two-Level Grid build (cell lists bit-exact) and 3D-DDA traversal (contacts bit-exact) through the C ABI,
against golden vectors produced by the reference's own TwoLevelGrid.h functions (tests/golden/tlg_*.npz) and, where
oracle/_ref/libclrt_ref.so is present, live against those functions on synthetic scenes."""
import os
import numpy as np
import pytest
from conftest import BUNDLED, ROOT, digest
from openclraytracer_b200 import api, scenes, layouts as L

pytestmark = pytest.mark.gpu
KEYS = ("top_cells", "top_ranges", "top_pairs", "leaf_ranges", "leaf_pairs")


def golden_tlg(name):
    with np.load(os.path.join(ROOT, "tests", "golden", f"tlg_{name}.npz")) as z:
        return {k: z[k] for k in z.files}


def build_gpu(ctx, scene_buf, top=2.0, leaf=2.0):
    scene = api.Scene(scene_buf); scene.loadToGPU(ctx)
    tlg = api.TwoLevelGridManager(ctx, scene, top, leaf); tlg.initialize(); tlg.initializeFrame(); tlg.construct()
    return scene, tlg


def check_contacts(got, want):
    hit = want["nd"][:, 3] > 0
    assert np.array_equal(got["nd"][:, 3] > 0, hit)
    assert np.array_equal(got["nd"][hit], want["nd"][hit])                       # normal + t bit-exact
    assert np.array_equal(got["materialIndex"][hit], want["materialIndex"][hit])
    assert np.array_equal(got["pixelIndex"], want["pixelIndex"])
    assert (got["nd"][~hit] == 0).all()


@pytest.mark.parametrize("name", BUNDLED)
def test_tlg_build_matches_reference_golden(ctx, name):
    g = golden_tlg(name)
    scene, tlg = build_gpu(ctx, scenes.load_bundled(name))
    assert tuple(int(x) for x in g["sizes"]) == tlg.sizes()
    e = tlg.export()
    assert e["grid"].tobytes() == g["grid"].tobytes()                            # GridData (64 B) bit-exact
    for k in KEYS:
        assert digest(e[k]) == str(g[k + "_sha"]), k
    # Testing/TwoLevelGridTest.h:133-156 (range/cell consistency) on the real pairs
    lp, lr = e["leaf_pairs"], e["leaf_ranges"]
    assert (lp["x"][1:] >= lp["x"][:-1]).all()
    nonempty = lr["y"] > lr["x"]
    assert (lr["y"][nonempty] - lr["x"][nonempty]).sum() == len(lp)
    tlg.free(); scene.free()


@pytest.mark.parametrize("name", BUNDLED)
def test_tlg_trace_matches_reference_golden(ctx, name):
    g = golden_tlg(name)
    scene, tlg = build_gpu(ctx, scenes.load_bundled(name))
    for k in range(2):
        tlg.generateContacts(g[f"cam{k}"])
        check_contacts(tlg.download_primary_contacts(), g[f"contacts{k}"])
    tlg.free(); scene.free()


@pytest.mark.parametrize("n,tpm,top,leaf", [(50, 16, 2.0, 2.0), (3000, 1000, 2.0, 2.0), (3000, 1000, 0.7, 5.0), (120_000, 21840, 2.0, 2.0)])
def test_tlg_soup_vs_reference_live(ctx, O, n, tpm, top, leaf):
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    buf = scenes.soup_scene(n, seed=100 + n, tris_per_mesh=tpm)
    scene, tlg = build_gpu(ctx, buf, top, leaf)
    ref = O.RefTlg(buf, top, leaf)
    assert ref.sizes() == tlg.sizes()
    e, r = tlg.export(), ref.export()
    assert e["grid"].tobytes() == r["grid"].tobytes()
    for k in KEYS:
        assert np.array_equal(e[k], r[k]), k
    cam = api.demo_camera(buf, 200, 120)
    tlg.generateContacts(cam)
    check_contacts(tlg.download_primary_contacts(), ref.trace_primary(cam))
    # generic-ray entry point
    nr = 200 * 120
    rays = api.Buffer(ctx, nr * 48); contacts = api.Buffer(ctx, nr * 32)
    api.generate_primary_rays(ctx, cam, rays.getCLMem())
    tlg.generateContacts(rays, contacts, nr)
    check_contacts(contacts.copyToHost(L.CONTACT, nr), ref.trace_rays(rays.copyToHost(L.RAY, nr)))
    rays.free(); contacts.free(); ref.free(); tlg.free(); scene.free()


def test_tlg_frame_and_stripes(ctx):
    """The frame entry points work over the grid as over the BVH: stripes compose to the one-rank frame."""
    buf = scenes.load_bundled("gardenlight")
    scene, tlg = build_gpu(ctx, buf)
    cam = api.demo_camera(buf, 160, 100)
    img, rays = tlg.render_frame_host(cam, shadows=True, want_rays=True)
    assert img[..., 3].any() and rays > 160 * 100
    cap, _ = api.frame_stripe_pixels(160, 100, 0, 2)
    stripes = api.Buffer(ctx, cap * 4 * 2); raster = api.Buffer(ctx, 160 * 100 * 4)
    for r in range(2):
        tlg.render_stripe(cam, True, r, 2, stripes.getCLMem() + r * cap * 4)
    api.frame_untile(ctx, stripes.getCLMem(), 160, 100, 2, raster.getCLMem())
    assert np.array_equal(raster.copyToHost(np.uint8).reshape(100, 160, 4), img)
    stripes.free(); raster.free(); tlg.free(); scene.free()


def test_tlg_rebuild_is_idempotent_and_densities_apply(ctx):
    buf = scenes.load_bundled("tenryuu")
    scene, tlg = build_gpu(ctx, buf)
    a = tlg.export()
    tlg.initializeFrame(); tlg.construct()
    b = tlg.export()
    for k in KEYS:
        assert np.array_equal(a[k], b[k])
    tlg.free()
    tlg2 = api.TwoLevelGridManager(ctx, scene); tlg2.setTopLevelDensity(0.5); tlg2.setLeafDensity(1.0)
    tlg2.initialize(); tlg2.initializeFrame(); tlg2.construct()
    assert tlg2.sizes()[0] < a["top_cells"].shape[0]
    tlg2.free(); scene.free()


def test_tlg_degenerate_scene_is_an_error(ctx):
    """A flat scene gives a zero top-level resolution; the reference divides by zero, here it is reported."""
    verts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], dtype=np.float32)
    buf = scenes.pack_scene([[0, 0, -5, 100]], [[(verts, np.array([0, 1, 2]), 0)]])
    scene = api.Scene(buf); scene.loadToGPU(ctx)
    tlg = api.TwoLevelGridManager(ctx, scene); tlg.initialize()
    with pytest.raises(api.ClrtError, match="degenerate"):
        tlg.initializeFrame()
    tlg.free(); scene.free()


@pytest.mark.parametrize("name", ["cube", "tenryuu", "clumsydragon", "soup"])
def test_corrected_grid_walk_returns_the_bvh_closest_hits(ctx, O, name):
    """Row f4 (KNOWN_ISSUES Q21): CLRT_TRACE_FAST on a grid selects the corrected two-level DDA (entry = max of the slab entries, the
    closest hit kept across cells, the walk ends when it lies in a finished cell).  Its contacts must be the TRUE closest hits —
    the oracle's BVH contacts — by the north-star tolerances, for camera rays and for generic rays; the default stays the literal walk."""
    from test_gpu_bvh import compare_fast
    buf = scenes.soup_scene(40_000, seed=11, tris_per_mesh=8192) if name == "soup" else scenes.load_bundled(name)
    scene, tlg = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    for cam in (api.demo_camera(buf, 320, 200), api.demo_camera(buf, 97, 61)):
        want = O.bvh_trace_primary(cam, b)
        tlg.set_trace_mode(api.TRACE_FAST)
        tlg.generateContacts(cam)
        got = tlg.download_primary_contacts()
        compare_fast(got, want, b["tris"])
        assert np.array_equal(got["materialIndex"][want["nd"][:, 3] > 0], want["materialIndex"][want["nd"][:, 3] > 0])
    n = 320 * 200
    cam = api.demo_camera(buf, 320, 200)
    rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_primary_rays(ctx, cam, rays.getCLMem())
    tlg.generateContacts(rays, contacts, n)
    compare_fast(contacts.copyToHost(L.CONTACT, n), O.bvh_trace_primary(cam, b), b["tris"])
    rays.free(); contacts.free(); tlg.free(); scene.free()
