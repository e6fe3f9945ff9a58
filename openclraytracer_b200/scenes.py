"""Scene buffers in the reference's packed v0.6 format + the deterministic synthetic scenes of SURVEY.md 8d.

Format (R/Include/CLData/SceneBufferParser.h; packer R/src/CLRayTracerEngine/Scene.cpp:231-444):
  SceneHeader(96) | Light[numLights](16) | Sphere[0] | Material[numMaterials](96) |
  per model: ModelHeader(64) | per sub-mesh: MeshHeader(48) | float3 verts (16 B each) | ushort indices
Sub-mesh blocks are NOT padded (Scene.cpp:80-86); indices are 16 bit (<= 65536 vertices per sub-mesh, Q13).
Bounding boxes start from (FLT_MAX.., FLT_MIN..) (Q4) so every max bound is >= FLT_MIN.
"""
import os
import numpy as np
from . import layouts as L

FLT_MAX = np.float32(3.402823466e+38)
FLT_MIN = np.float32(1.175494351e-38)
GOLDEN_SEED = 20161116


def default_material():
    m = np.zeros((), dtype=L.MATERIAL)
    m["ambient"][:3] = 0.1
    m["diffuse"][:3] = (0.7, 0.7, 0.7)
    m["specular"][:3] = (0.3, 0.3, 0.3)
    m["shininess"] = 8.0
    m["ior"] = 1.0
    m["dissolve"] = 1.0
    m["illum"] = 2.0
    return m


def pack_scene(lights, models, materials=None):
    """lights: (nl,4) float32 (x,y,z,energy).  models: list of models; a model is a list of sub-meshes
    (verts (nv,3) float32, indices (ni,) int, material_index).  Returns the packed buffer as uint8 array."""
    lights = np.asarray(lights, dtype=np.float32).reshape(-1, 4)
    if materials is None:
        materials = np.array([default_material()], dtype=L.MATERIAL)
    materials = np.asarray(materials, dtype=L.MATERIAL).reshape(-1)
    model_sizes, total_tris = [], 0
    for model in models:
        sz = L.MODEL_HEADER.itemsize
        for verts, idx, _mat in model:
            verts = np.asarray(verts); idx = np.asarray(idx)
            if len(verts) > 65536:
                raise ValueError("a sub-mesh holds at most 65536 vertices (16-bit indices, SceneBufferParser.h:160)")
            if len(idx) % 3:
                raise ValueError("index count must be a multiple of 3")
            sz += L.MESH_HEADER.itemsize + 16 * len(verts) + 2 * len(idx)
            total_tris += len(idx) // 3
        model_sizes.append(sz)
    model_bytes = int(sum(model_sizes))
    total = L.SCENE_HEADER.itemsize + 16 * len(lights) + 96 * len(materials) + model_bytes
    buf = np.zeros(total, dtype=np.uint8)
    hdr = buf[:96].view(L.SCENE_HEADER)[0]
    hdr["totalDataSize"] = total
    hdr["numberOfPrimitives"] = len(lights) + len(models)
    hdr["numberOfLights"] = len(lights)
    hdr["numberOfSpheres"] = 0
    hdr["numberOfMaterials"] = len(materials)
    hdr["modelBufferSize"] = model_bytes
    hdr["numberOfModels"] = len(models)
    hdr["totalNumberOfTriangles"] = total_tris
    off = 96
    buf[off:off + 16 * len(lights)] = lights.view(np.uint8).reshape(-1)
    off += 16 * len(lights)
    buf[off:off + 96 * len(materials)] = materials.view(np.uint8).reshape(-1)
    off += 96 * len(materials)
    smin = np.array([FLT_MAX] * 3, dtype=np.float32)
    smax = np.array([FLT_MIN] * 3, dtype=np.float32)
    for model, msz in zip(models, model_sizes):
        mh = buf[off:off + 64].view(L.MODEL_HEADER)[0]
        mh["dataSize"] = msz
        mh["numberOfSubmeshes"] = len(model)
        mmin = np.array([FLT_MAX] * 3, dtype=np.float32)
        mmax = np.array([FLT_MIN] * 3, dtype=np.float32)
        moff = off + 64
        mtris = 0
        for verts, idx, mat in model:
            verts = np.ascontiguousarray(verts, dtype=np.float32).reshape(-1, 3)
            idx = np.ascontiguousarray(idx).reshape(-1)
            sh = np.zeros((), dtype=L.MESH_HEADER)
            sh["dataSize"] = 48 + 16 * len(verts) + 2 * len(idx)
            sh["numberOfTriangles"] = len(idx) // 3
            sh["numberOfVertices"] = len(verts)
            sh["numberOfIndices"] = len(idx)
            sh["materialIndex"] = mat
            buf[moff:moff + 48] = np.frombuffer(sh.tobytes(), dtype=np.uint8)
            v4 = np.zeros((len(verts), 4), dtype=np.float32)
            v4[:, :3] = verts
            buf[moff + 48:moff + 48 + 16 * len(verts)] = v4.view(np.uint8).reshape(-1)
            i16 = idx.astype(np.uint16)                      # Scene.cpp:429-430 truncates the same way
            buf[moff + 48 + 16 * len(verts):moff + 48 + 16 * len(verts) + 2 * len(idx)] = i16.view(np.uint8)
            if len(verts):
                mmin = np.minimum(mmin, verts.min(axis=0))
                mmax = np.maximum(mmax, verts.max(axis=0))
            mtris += len(idx) // 3
            moff += int(sh["dataSize"])
        mh["numberOfTriangles"] = mtris
        mh["boundingBox"][0, :3] = mmin
        mh["boundingBox"][1, :3] = mmax
        smin = np.minimum(smin, mmin)
        smax = np.maximum(smax, mmax)
        off += msz
    hdr["modelsBoundingBox"][0, :3] = smin
    hdr["modelsBoundingBox"][1, :3] = smax
    return buf


# ---- deterministic generators -----------------------------------------------------------------
def splitmix64(x):
    x = np.asarray(x, dtype=np.uint64)
    with np.errstate(over="ignore"):
        z = x + np.uint64(0x9E3779B97F4A7C15)
        z = (z ^ (z >> np.uint64(30))) * np.uint64(0xBF58476D1CE4E5B9)
        z = (z ^ (z >> np.uint64(27))) * np.uint64(0x94D049BB133111EB)
        return z ^ (z >> np.uint64(31))


def uniform01(seed, first, count):
    """float32 in [0,1): (splitmix64(seed*0x9E3779B97F4A7C15 + counter) >> 40) * 2^-24"""
    with np.errstate(over="ignore"):
        ctr = np.uint64(seed) * np.uint64(0x9E3779B97F4A7C15) + np.arange(first, first + count, dtype=np.uint64)
    return ((splitmix64(ctr) >> np.uint64(40)).astype(np.float32) * np.float32(2.0 ** -24)).astype(np.float32)


def soup_triangles(n, seed=GOLDEN_SEED):
    """n random triangles: centroid uniform in [0,1)^3, two edge vectors uniform in [-s,s]^3, s = 0.5*n^(-1/3).
    Returns (n,3,3) float32 vertices."""
    s = np.float32(0.5 * float(n) ** (-1.0 / 3.0))
    u = uniform01(seed, 0, 9 * n).reshape(n, 9)
    c = u[:, 0:3]
    e1 = (u[:, 3:6] * np.float32(2.0) - np.float32(1.0)) * s
    e2 = (u[:, 6:9] * np.float32(2.0) - np.float32(1.0)) * s
    v0 = c - (e1 + e2) / np.float32(3.0)
    tri = np.stack([v0, v0 + e1, v0 + e2], axis=1).astype(np.float32)
    return tri


def soup_scene(n, seed=GOLDEN_SEED, tris_per_mesh=21840):
    """Soup-N as one model of ceil(n/21840) sub-meshes with unshared vertices (21840*3 = 65520 <= 2^16; 21840 % 8 == 0
    keeps every mesh block 16-byte aligned), one material, one light at (0.5, 10, -3.3) energy 1000."""
    tri = soup_triangles(n, seed)
    model = []
    for first in range(0, n, tris_per_mesh):
        t = tri[first:first + tris_per_mesh]
        verts = t.reshape(-1, 3)
        idx = np.arange(len(verts), dtype=np.uint32)
        model.append((verts, idx, 0))
    lights = np.array([[0.5, 10.0, -3.3, 1000.0]], dtype=np.float32)
    return pack_scene(lights, [model])


def sphere_field_scene(nspheres=2000, seg=50, seed=GOLDEN_SEED, radius=0.35):
    """Tessellated-sphere field: `nspheres` lat-long spheres (seg x seg quads -> 2*seg*seg triangles, (seg+1)^2 shared
    vertices, one sub-mesh each), centres on a jittered 20x10x10 lattice in [0,20]x[0,10]x[0,10]."""
    nx, ny, nz = 20, 10, 10
    if nspheres > nx * ny * nz:
        raise ValueError("at most 2000 spheres")
    jit = uniform01(seed, 0, 3 * nspheres).reshape(nspheres, 3)
    ids = np.arange(nspheres)
    centres = np.stack([ids % nx, (ids // nx) % ny, ids // (nx * ny)], axis=1).astype(np.float32) + np.float32(0.5)
    centres = centres + (jit - np.float32(0.5)) * np.float32(0.25)
    th = np.linspace(0.0, np.pi, seg + 1, dtype=np.float64)
    ph = np.linspace(0.0, 2.0 * np.pi, seg + 1, dtype=np.float64)
    T, P = np.meshgrid(th, ph, indexing="ij")
    unit = np.stack([np.sin(T) * np.cos(P), np.cos(T), np.sin(T) * np.sin(P)], axis=-1).reshape(-1, 3).astype(np.float32)
    i, j = np.meshgrid(np.arange(seg), np.arange(seg), indexing="ij")
    a = (i * (seg + 1) + j).reshape(-1); b = a + 1; c = a + (seg + 1); d = c + 1
    idx = np.stack([a, c, b, b, c, d], axis=1).reshape(-1).astype(np.uint32)
    model = [(unit * np.float32(radius) + centres[s], idx, 0) for s in range(nspheres)]
    lights = np.array([[10.0, 30.0, -10.0, 1000.0]], dtype=np.float32)
    return pack_scene(lights, [model])


def cube_scene():
    """A 12-triangle unit cube as 6 two-triangle sub-meshes (mis-aligned blocks like the bundled Cube, Q14)."""
    model = []
    for axis in range(3):
        for side in (0.0, 1.0):
            u, v = [a for a in range(3) if a != axis]
            q = np.zeros((4, 3), dtype=np.float32)
            q[:, axis] = side
            q[1, u] = 1; q[2, u] = 1; q[2, v] = 1; q[3, v] = 1
            model.append((q - np.float32(0.5), np.array([0, 1, 2, 0, 2, 3], dtype=np.uint32), 0))
    lights = np.array([[0.0, 10.0, -330.0, 1000.0]], dtype=np.float32)
    return pack_scene(lights, [model])


# ---- bundled meshes: packed by the reference's own loader, stored as fixtures ---------------------------------
GOLDEN_DIR = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def load_bundled(name):
    """Packed scene buffer of a bundled TestModels mesh (cube, gardenlight, tenryuu, manfigure, clumsydragon) from
    tests/golden/scenes/<name>.npz — produced by tests/golden/make_golden.py with the reference's tinyobj + packer."""
    path = os.path.join(GOLDEN_DIR, "scenes", f"{name}.npz")
    with np.load(path) as z:
        return np.ascontiguousarray(z["scene"])


def scene_header(buf):
    return np.frombuffer(buf[:96].tobytes(), dtype=L.SCENE_HEADER)[0]


def triangle_count(buf):
    return int(scene_header(buf)["totalNumberOfTriangles"])
