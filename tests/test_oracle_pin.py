"""Pins the oracle (oracle/clrt_oracle.cpp, our CPU restatement) to the REFERENCE:
 (1) against the committed golden vectors, which were produced by the reference's own headers + tinyobj compiled in
     place (tests/golden/make_golden.py -> oracle/_ref/libclrt_ref.so);
 (2) live against oracle/_ref/libclrt_ref.so where it exists (build container; it also travels to the GPU box).
The reference itself holds no golden vectors or known-answer tests."""
import numpy as np
import pytest
from conftest import BUNDLED, masked_nodes, digest
from openclraytracer_b200 import scenes, layouts as L


@pytest.mark.parametrize("name", BUNDLED)
def test_oracle_build_matches_golden(O, golden, name):
    g = golden(name)
    scene = scenes.load_bundled(name)
    b = O.bvh_build(scene)
    n = b["n"]
    assert n == int(g["n"])
    tris, tim = b["tris"], b["tim"]
    _, unsorted = O.bvh_morton(scene, tris, tim)
    assert digest(unsorted) == str(g["unsorted_sha"])        # Morton codes bit-exact
    assert digest(b["pairs"]) == str(g["pairs_sha"])          # canonical sorted order
    assert digest(masked_nodes(b["nodes"], n)) == str(g["nodes_sha"])   # topology + boxes, bit-exact
    if "nodes" in g:
        assert np.array_equal(masked_nodes(b["nodes"], n), g["nodes"])


@pytest.mark.parametrize("name", BUNDLED)
def test_oracle_trace_and_shade_match_golden(O, golden, name):
    g = golden(name)
    scene = scenes.load_bundled(name)
    b = O.bvh_build(scene)
    for k in range(2):
        cam = g[f"cam{k}"]
        want = g[f"contacts{k}"]
        got = O.bvh_trace_primary(cam, b)
        hit = want["nd"][:, 3] > 0
        assert hit.any()
        assert np.array_equal(got["nd"][:, 3] > 0, hit)
        assert np.array_equal(got["nd"][hit], want["nd"][hit])                  # normal + t, bit-exact
        assert np.array_equal(got["materialIndex"][hit], want["materialIndex"][hit])
        assert np.array_equal(got["pixelIndex"], want["pixelIndex"])
        _, f32 = O.shade(cam, scene, got, want_f32=True)
        assert np.array_equal(f32[hit], g[f"colors{k}"][hit])                   # host Blinn-Phong, bit-exact
        assert not f32[~hit].any()


def test_demo_camera_matches_golden(O, golden):
    from openclraytracer_b200 import api
    for name in BUNDLED:
        g = golden(name)
        scene = scenes.load_bundled(name)
        cam = O.demo_camera(scene, 96, 64)
        assert cam.tobytes() == g["cam0"].tobytes()
        assert api.demo_camera(scene, 96, 64).tobytes() == g["cam0"].tobytes()   # host-side mirror (api.py)


def test_ray_generation_matches_reference(O, golden):
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    g = golden("tenryuu")
    for k in range(2):
        a = O.generate_primary_rays(g[f"cam{k}"])
        b = O.ref_generate_rays(g[f"cam{k}"])
        assert np.array_equal(a["origin"][:, :3], b["origin"][:, :3])
        assert np.array_equal(a["direction"][:, :3], b["direction"][:, :3])


def test_struct_sizes_match_reference(O):
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    s = O.ref_struct_sizes()
    names = ["BVH_NODE", "AABB", "CONTACT", "CAMERA", None, "SCENE_HEADER", "MODEL_HEADER", "MESH_HEADER", "MATERIAL", "LIGHT",
             "GRID_DATA", "TOP_CELL", "FLOAT4", "UINT2"]
    for size, name in zip(s, names):
        if name:
            assert size == getattr(L, name).itemsize, name
    assert s[4] == 48            # Matrix4


@pytest.mark.parametrize("n", [1, 2, 3, 7, 64, 1000])
def test_oracle_vs_reference_live_on_soup(O, n):
    """Synthetic scenes (our packer) through the reference's own headers vs the restatement: codes, topology, boxes,
    contacts.  Covers N=1 (no inner node), N=2, odd sizes."""
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    scene = scenes.soup_scene(n, seed=7 + n, tris_per_mesh=24)
    b = O.bvh_build(scene)
    r = O.ref_bvh_build(scene)
    assert np.array_equal(b["pairs"], r["pairs"])
    assert np.array_equal(masked_nodes(b["nodes"], n)[: 2 * n - 1], masked_nodes(r["nodes"], n)[: 2 * n - 1])
    if n >= 2:
        cam = O.demo_camera(scene, 48, 32)
        a = O.bvh_trace_primary(cam, b)
        c = O.ref_trace_primary(cam, r["nodes"], n, scene)
        hit = c["nd"][:, 3] > 0
        assert np.array_equal(a["nd"][:, 3] > 0, hit)
        assert np.array_equal(a["nd"][hit], c["nd"][hit])


def test_literal_mode_differs_only_in_last_inner_node(O):
    """Q1: the reference's guard skips inner node N-2; LITERAL leaves slot 2N-2 unbuilt, all other slots identical."""
    scene = scenes.soup_scene(100, seed=3, tris_per_mesh=48)
    tris, tim = O.flatten(scene)
    nodes, pairs = O.bvh_morton(scene, tris, tim)
    pairs = O.sort_pairs_stable(pairs)
    import ctypes as C
    n = 100
    fixed, lit = nodes.copy(), nodes.copy()
    O.oracle().oracle_bvh_radix_tree(fixed.ctypes.data_as(C.c_void_p), pairs.ctypes.data_as(C.c_void_p), C.c_uint32(n), C.c_int(0))
    O.oracle().oracle_bvh_radix_tree(lit.ctypes.data_as(C.c_void_p), pairs.ctypes.data_as(C.c_void_p), C.c_uint32(n), C.c_int(1))
    assert np.array_equal(fixed["data"][n:2 * n - 2, 1:3], lit["data"][n:2 * n - 2, 1:3])
    assert not np.array_equal(fixed["data"][2 * n - 2, 1:3], lit["data"][2 * n - 2, 1:3])


@pytest.mark.parametrize("kind", __import__("edge_scenes").KINDS)
def test_oracle_vs_reference_live_on_edge_scenes(O, kind):
    """Zero extent on an axis, all-negative boxes, identical triangles, zero-area triangles, collapsed axes, 1e9 dynamic range
    (tests/edge_scenes.py): whatever the reference's headers produce — inf/NaN normalisation, clamped cells, chains — the
    restatement produces the same bits: pairs, nodes, contacts."""
    import edge_scenes
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    scene = edge_scenes.edge_scene(kind)
    b = O.bvh_build(scene)
    r = O.ref_bvh_build(scene)
    n = len(b["tris"])
    assert np.array_equal(b["pairs"], r["pairs"])
    assert np.array_equal(masked_nodes(b["nodes"], n)[: 2 * n - 1], masked_nodes(r["nodes"], n)[: 2 * n - 1])
    cam = edge_scenes.edge_camera(kind, 64, 40)
    a = O.bvh_trace_primary(cam, b)
    c = O.ref_trace_primary(cam, r["nodes"], n, scene)
    hit = c["nd"][:, 3] > 0
    assert hit.sum() > 10, "the camera must see the geometry"
    assert np.array_equal(a["nd"][:, 3] > 0, hit)
    assert np.array_equal(a["nd"][hit], c["nd"][hit])


@pytest.mark.parametrize("seed", range(6))
def test_shading_restatement_vs_reference_live_with_many_lights_and_materials(O, seed):
    """cpuShadeBlinnPhong (Shading.h:74-166) through the reference's own header vs the restatement, on scenes the bundled models
    do not have: 0-4 lights (some inside the geometry, one at the camera), 1-5 materials with random coefficients, shininess from
    0 to 200, zero and >1 colour components (clamping), sub-meshes cycling through the materials.  Bit-exact float4 colours."""
    if not O.have_ref():
        pytest.skip("oracle/_ref not built")
    rng = np.random.default_rng(1000 + seed)
    n = 400
    tri = scenes.soup_triangles(n, 50 + seed)
    nm = int(rng.integers(1, 6))
    mats = np.zeros(nm, dtype=L.MATERIAL)
    for m in mats:
        for key in ("ambient", "diffuse", "specular", "transmittance", "emission"):
            m[key][:3] = rng.uniform(0.0, 1.6, 3).astype(np.float32) * (rng.random(3) > 0.15)
        m["shininess"] = np.float32(rng.choice([0.0, 1.0, 8.0, 37.5, 200.0]))
        m["ior"], m["dissolve"], m["illum"] = 1.0, 1.0, 2.0
    model = []
    for i, first in enumerate(range(0, n, 64)):
        verts = tri[first:first + 64].reshape(-1, 3)
        model.append((verts, np.arange(len(verts), dtype=np.uint32), i % nm))
    nl = int(rng.integers(0, 5))
    lights = np.zeros((nl, 4), dtype=np.float32)
    for i in range(nl):
        lights[i, :3] = rng.uniform(-0.5, 1.5, 3)
        lights[i, 3] = np.float32(rng.choice([0.0, 1.0, 50.0, 1000.0]))
    scene = scenes.pack_scene(lights, [model], materials=mats)
    cam = O.demo_camera(scene, 80, 60)
    if nl:                                                       # a light exactly at the eye
        lights[0, :3] = cam["viewTransform"][[3, 7, 11]]
        scene = scenes.pack_scene(lights, [model], materials=mats)
    b = O.bvh_build(scene)
    c = O.bvh_trace_primary(cam, b)
    hit = c["nd"][:, 3] > 0
    assert hit.sum() > 200 and len(np.unique(c["materialIndex"][hit])) == nm
    _, f32 = O.shade(cam, scene, c, want_f32=True)
    ref = O.ref_shade(cam, c, scene)
    assert np.array_equal(f32[hit].view(np.uint32), ref[hit].view(np.uint32))
    assert np.isfinite(f32[hit]).all()
