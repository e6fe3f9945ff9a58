"""GPU: incoherent secondary rays (BASELINE.json configs[3]) — one cosine-weighted diffuse bounce per primary hit,
generated on the device and traced through the reference's generic-ray entry point (generateContacts(rays, contacts, n)).
Generator vs a numpy restatement (1e-5: sincos / sqrt differ in the last ulps between CUDA and libm); traced contacts vs the oracle on
the SAME rays (bit-exact in REFERENCE_ORDER mode, north-star tolerances in FAST mode); Spheres-10M at 4K through size-independent
properties."""
import numpy as np
import pytest
from openclraytracer_b200 import api, scenes, layouts as L
from test_gpu_bvh import build_gpu, compare_fast

pytestmark = pytest.mark.gpu
SEED = 20161116
M64 = (1 << 64) - 1


def splitmix64(x):
    x = (x + 0x9E3779B97F4A7C15) & M64
    x = ((x ^ (x >> 30)) * 0xBF58476D1CE4E5B9) & M64
    x = ((x ^ (x >> 27)) * 0x94D049BB133111EB) & M64
    return x ^ (x >> 31)


def bounce_reference(cam, contacts, prim_rays, seed):
    """numpy (float64 inside, compared at 1e-5) restatement of bounce_rays_kernel."""
    hit = contacts["nd"][:, 3] > 0
    o = prim_rays["origin"][:, :3].astype(np.float64); d = prim_rays["direction"][:, :3].astype(np.float64)
    t = contacts["nd"][:, 3].astype(np.float64)[:, None]
    n = contacts["nd"][:, :3].astype(np.float64).copy()
    flip = (n * d).sum(axis=1) > 0
    n[flip] *= -1
    P = o + d * t
    pix = contacts["pixelIndex"].astype(object)
    u = np.array([splitmix64((seed * 0x9E3779B97F4A7C15 + int(p)) & M64) for p in pix], dtype=object)
    u1 = np.array([int(v) >> 40 for v in u], dtype=np.float64) / 16777216.0
    u2 = np.array([int(v) & 0xFFFFFF for v in u], dtype=np.float64) / 16777216.0
    r = np.sqrt(u1); phi = 2 * np.pi * u2
    dx, dy, dz = r * np.cos(phi), r * np.sin(phi), np.sqrt(np.maximum(0.0, 1.0 - u1))
    ax = np.where((np.abs(n[:, 0]) < 0.5)[:, None], np.array([1.0, 0, 0]), np.array([0, 1.0, 0]))
    t1 = np.cross(ax, n); t1 /= np.linalg.norm(t1, axis=1, keepdims=True) + 1e-300
    t2 = np.cross(n, t1)
    dirs = t1 * dx[:, None] + t2 * dy[:, None] + n * dz[:, None]
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True) + 1e-300
    org = P + n * (1e-4 * t)
    return hit, org, dirs, n


@pytest.mark.parametrize("name", ["tenryuu", "soup"])
def test_bounce_rays_and_their_contacts(ctx, O, name):
    buf = scenes.soup_scene(20_000, seed=3, tris_per_mesh=4096) if name == "soup" else scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    W, H = 192, 108
    cam = api.demo_camera(buf, W, H)
    n = W * H
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    prim = bvh.download_primary_contacts()
    prays = api.Buffer(ctx, n * 48); rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_primary_rays(ctx, cam, prays.getCLMem())
    api.generate_bounce_rays(ctx, cam, bvh.getPrimaryContacts(), n, SEED, rays.getCLMem())
    r = rays.copyToHost(L.RAY, n)
    hit, org, dirs, nrm = bounce_reference(cam, prim, prays.copyToHost(L.RAY, n), SEED)
    assert hit.sum() > 500
    assert (r["idx"][~hit] == 0xFFFFFFFF).all() and (r["direction"][~hit] == 0).all()
    assert np.array_equal(r["idx"][hit], prim["pixelIndex"][hit])
    scale = np.abs(org[hit]).max()
    assert np.abs(r["direction"][hit, :3] - dirs[hit]).max() < 2e-5
    assert np.abs(r["origin"][hit, :3] - org[hit]).max() < 2e-5 * max(scale, 1.0)
    assert np.abs(np.linalg.norm(r["direction"][hit, :3].astype(np.float64), axis=1) - 1).max() < 1e-6
    cosn = (r["direction"][hit, :3].astype(np.float64) * nrm[hit]).sum(axis=1)
    assert (cosn > -1e-6).all() and 0.55 < cosn.mean() < 0.78          # cosine-weighted: E[cos] = 2/3
    want = O.bvh_trace_rays(r, b)
    for mode in (api.TRACE_REFERENCE_ORDER, api.TRACE_FAST):
        bvh.set_trace_mode(mode)
        bvh.generateContacts(rays, contacts, n)
        got = contacts.copyToHost(L.CONTACT, n)
        assert (got["nd"][~hit, 3] == 0).all()
        if mode == api.TRACE_REFERENCE_ORDER:
            h2 = want["nd"][:, 3] > 0
            assert np.array_equal(got["nd"][:, 3] > 0, h2) and np.array_equal(got["nd"][h2], want["nd"][h2])
            assert np.array_equal(got["pad"][h2, 0], want["pad"][h2, 0])
        else:
            compare_fast(got, want, b["tris"])
    for x in (prays, rays, contacts):
        x.free()
    bvh.free(); scene.free()


def moller_trumbore_t(tri, o, d):
    v0, v1, v2 = tri[:, 0].astype(np.float64), tri[:, 1].astype(np.float64), tri[:, 2].astype(np.float64)
    e1, e2 = v1 - v0, v2 - v0
    p = np.cross(d, e2); det = (e1 * p).sum(axis=1)
    tv = o - v0; u = (tv * p).sum(axis=1) / det
    q = np.cross(tv, e1); v = (d * q).sum(axis=1) / det
    return (e2 * q).sum(axis=1) / det, u, v


def test_spheres_10m_incoherent_bounce_at_4k(ctx):
    """configs[3] at full size: 10 M triangles (2000 tessellated spheres), 3840x2160 primary rays, one diffuse bounce per hit
    through the generic-ray entry.  Properties that do not need the oracle: every reported hit re-intersects its own triangle at
    the reported t (float64 Moeller-Trumbore on a sample), misses of the primary pass stay misses, FAST agrees with
    REFERENCE_ORDER on a strided sample of the same bounce rays."""
    buf = scenes.sphere_field_scene()
    scene, bvh = build_gpu(ctx, buf)
    assert scene.triangle_count() == 10_000_000
    W, H = 3840, 2160
    n = W * H
    cam = api.demo_camera(buf, W, H)
    bvh.set_trace_mode(api.TRACE_FAST)
    bvh.generateContacts(cam)
    prim = bvh.download_primary_contacts()
    hit = prim["nd"][:, 3] > 0
    assert hit.mean() > 0.3
    rays = api.Buffer(ctx, n * 48); contacts = api.Buffer(ctx, n * 32)
    api.generate_bounce_rays(ctx, cam, bvh.getPrimaryContacts(), n, SEED, rays.getCLMem())
    bvh.generateContacts(rays, contacts, n)
    got = contacts.copyToHost(L.CONTACT, n)
    assert (got["nd"][~hit, 3] == 0).all()
    h2 = got["nd"][:, 3] > 0
    assert 0.05 < h2[hit].mean() < 0.95                              # some bounces hit a neighbouring sphere, some leave the field
    r = rays.copyToHost(L.RAY, n)
    tris = scene.export_triangles()
    sample = np.flatnonzero(h2)[:: max(1, int(h2.sum()) // 200_000)]
    tri = tris[got["pad"][sample, 0]]
    tri = np.stack([tri["v0"][:, :3], tri["v1"][:, :3], tri["v2"][:, :3]], axis=1) if tri.dtype.names else tri.reshape(-1, 3, 4)[:, :, :3]
    t, u, v = moller_trumbore_t(tri, r["origin"][sample, :3].astype(np.float64), r["direction"][sample, :3].astype(np.float64))
    tg = got["nd"][sample, 3].astype(np.float64)
    assert (np.abs(t - tg) <= 1e-4 * np.abs(tg) + 1e-6).all()
    assert (u > -1e-4).all() and (v > -1e-4).all() and (u + v < 1 + 1e-4).all()
    # FAST vs REFERENCE_ORDER on a strided subset of the same rays (the unpruned reference order is slow on 10 M triangles)
    sub = np.arange(0, n, 257, dtype=np.int64)
    rs = api.Buffer.from_array(ctx, np.ascontiguousarray(r[sub])); cs = api.Buffer(ctx, len(sub) * 32)
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(rs, cs, len(sub))
    ref = cs.copyToHost(L.CONTACT, len(sub))
    compare_fast(got[sub], ref, None)
    for x in (rays, contacts, rs, cs):
        x.free()
    bvh.free(); scene.free()


def reflect_reference(o, d, contacts):
    """numpy (float64) restatement of secondary_rays_kernel<REFLECT>: rays (o, d) with their contacts -> mirrored rays."""
    t = contacts["nd"][:, 3].astype(np.float64)[:, None]
    n = contacts["nd"][:, :3].astype(np.float64).copy()
    o = o.astype(np.float64); d = d.astype(np.float64)
    n[(n * d).sum(axis=1) > 0] *= -1
    P = o + d * t
    r = d - n * (2.0 * (n * d).sum(axis=1))[:, None]
    r /= np.linalg.norm(r, axis=1, keepdims=True) + 1e-300
    return P + n * (1e-4 * t), r, n


@pytest.mark.parametrize("name", ["tenryuu", "soup"])
def test_reflection_depth_loop(ctx, O, name):
    """The depth loop of the reference's TODO 2.2: rays_1 from the primary contacts, rays_k+1 from the contacts of rays_k (the pixel
    travels in idx), three levels deep.  Every level: generator vs numpy, its contacts vs the oracle on the same rays, dead rays stay
    dead, and writing the new rays over the old ones (aliasing) gives the same bytes."""
    buf = scenes.soup_scene(20_000, seed=5, tris_per_mesh=4096) if name == "soup" else scenes.load_bundled(name)
    scene, bvh = build_gpu(ctx, buf)
    b = O.bvh_build(buf)
    W, H = 160, 90
    cam = api.demo_camera(buf, W, H)
    n = W * H
    bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
    bvh.generateContacts(cam)
    prev_c = bvh.download_primary_contacts()
    prays = api.Buffer(ctx, n * 48)
    api.generate_primary_rays(ctx, cam, prays.getCLMem())
    prev_r = prays.copyToHost(L.RAY, n)
    pixel = prev_c["pixelIndex"].copy()
    bufs = [api.Buffer(ctx, n * 48) for _ in range(2)]
    contacts = api.Buffer(ctx, n * 32)
    alive_counts = []
    for level in range(1, 4):
        out = bufs[level & 1]
        if level == 1:
            api.generate_secondary_rays(ctx, api.RAYS_REFLECT, cam, None, bvh.getPrimaryContacts(), n, 0, out.getCLMem())
        else:
            api.generate_secondary_rays(ctx, api.RAYS_REFLECT, None, bufs[(level - 1) & 1].getCLMem(), contacts.getCLMem(), n, 0, out.getCLMem())
        r = out.copyToHost(L.RAY, n)
        alive = (prev_c["nd"][:, 3] > 0) & (prev_r["idx"] != 0xFFFFFFFF)
        alive_counts.append(int(alive.sum()))
        org, dirs, nrm = reflect_reference(prev_r["origin"][:, :3], prev_r["direction"][:, :3], prev_c)
        assert (r["idx"][~alive] == 0xFFFFFFFF).all() and (r["direction"][~alive] == 0).all()
        assert np.array_equal(r["idx"][alive], pixel[alive])
        if alive.any():
            scale = max(np.abs(org[alive]).max(), 1.0)
            assert np.abs(r["direction"][alive, :3] - dirs[alive]).max() < 2e-5
            assert np.abs(r["origin"][alive, :3] - org[alive]).max() < 2e-5 * scale
            din = prev_r["direction"][alive, :3].astype(np.float64)
            # mirror law: the normal component flips, the tangential one stays
            assert np.abs((r["direction"][alive, :3] * nrm[alive]).sum(axis=1) + (din * nrm[alive]).sum(axis=1)).max() < 2e-5
        if level > 1:                                   # aliasing: new rays written over the rays they were made from
            tmp = api.Buffer.from_array(ctx, np.ascontiguousarray(prev_r))
            api.generate_secondary_rays(ctx, api.RAYS_REFLECT, None, tmp.getCLMem(), contacts.getCLMem(), n, 0, tmp.getCLMem())
            assert tmp.copyToHost(L.RAY, n).tobytes() == r.tobytes()
            tmp.free()
        want = O.bvh_trace_rays(r, b)
        bvh.set_trace_mode(api.TRACE_FAST)
        bvh.generateContacts(out, contacts, n)
        compare_fast(contacts.copyToHost(L.CONTACT, n), want, b["tris"])
        bvh.set_trace_mode(api.TRACE_REFERENCE_ORDER)
        bvh.generateContacts(out, contacts, n)
        got = contacts.copyToHost(L.CONTACT, n)
        h2 = want["nd"][:, 3] > 0
        assert np.array_equal(got["nd"][:, 3] > 0, h2) and np.array_equal(got["nd"][h2], want["nd"][h2])
        assert np.array_equal(got["pad"][h2, 0], want["pad"][h2, 0])
        assert (got["nd"][~alive, 3] == 0).all()
        prev_c, prev_r = got, r
    assert alive_counts[0] > 500 and alive_counts[0] >= alive_counts[1] >= alive_counts[2]
    if name == "soup":
        assert alive_counts[2] > 50                     # a soup keeps mirrors bouncing
    for x in (prays, contacts, *bufs):
        x.free()
    bvh.free(); scene.free()


@pytest.mark.parametrize("depth", ["1", "3"])
def test_bounce_renderer_frames_in_flight_equal_the_four_calls(ctx, depth, monkeypatch):
    """frame.BounceRenderer (what bench.py --workload spheres10m times: clrt_render_bounce_stripe_on on frame slots, several frames in
    flight on several streams, untile) against the four public calls made one after the other on the context's stream — the path the
    tests above check against the oracle.  More frames than slots; every raster must be identical."""
    from openclraytracer_b200 import frame
    monkeypatch.setenv("CLRT_BOUNCE_DEPTH", depth)
    buf = scenes.soup_scene(60_000, seed=4)
    W, H = 448, 264                                                     # not a multiple of the 32x8 tiles in y
    r = frame.BounceRenderer(buf, W, H, 0, 1, 0)
    assert r.DEPTH == int(depth)
    r.build()
    poses = list(range(2 * r.DEPTH + 1))
    got = []
    for p in poses:
        got.append(r.render_host(frame.orbit_camera(buf, W, H, p)))
        if (p + 1) % r.DEPTH == 0:                                       # host rasters are about to be reused: collect
            r.finish()
            got = [g.numpy().copy() if hasattr(g, "numpy") else g for g in got]
    r.finish()
    got = [g.numpy().copy() if hasattr(g, "numpy") else g for g in got]
    r.close()
    # the same frames by single calls
    scene, bvh = build_gpu(ctx, buf)
    cap, tiles = api.frame_stripe_pixels(W, H, 0, 1)
    n = tiles * frame.TILE_PIXELS
    contacts = api.Buffer(ctx, cap * 32); rays = api.Buffer(ctx, cap * 48); bnc = api.Buffer(ctx, cap * 32)
    stripe = api.Buffer(ctx, cap * 4); raster = api.Buffer(ctx, W * H * 4)
    bvh.set_trace_mode(api.TRACE_FAST)
    for p, g in zip(poses, got):
        cam = frame.orbit_camera(buf, W, H, p)
        bvh.trace_primary_stripe(cam, 0, 1, contacts.getCLMem())
        api.generate_bounce_rays(ctx, cam, contacts.getCLMem(), n, SEED, rays.getCLMem())
        bvh.generateContacts(rays.getCLMem(), bnc.getCLMem(), n)
        api.shade_bounce(ctx, scene, cam, contacts.getCLMem(), n, bnc.getCLMem(), stripe.getCLMem())
        api.frame_untile(ctx, stripe.getCLMem(), W, H, 1, raster.getCLMem())
        want = raster.copyToHost(np.int32, W * H)
        assert np.array_equal(np.asarray(g).reshape(-1), want), f"pose {p}"
        assert (want != 0).mean() > 0.2
    for b in (contacts, rays, bnc, stripe, raster):
        b.free()
    bvh.free(); scene.free()
