"""Edge-case scenes for the BVH path: what the reference's arithmetic does on them is whatever its headers do —
the restatement, the reference headers compiled in place and the CUDA path must agree bit for bit."""
import numpy as np
from openclraytracer_b200 import scenes

KINDS = ["planar_z", "negative", "duplicates", "degenerate", "line_x", "mixed_sign", "huge_and_tiny"]


def edge_triangles(kind, n=600, seed=11):
    tri = scenes.soup_triangles(n, seed).copy()
    if kind == "planar_z":                      # zero extent on one axis: c/(max-min) divides by zero (Q6)
        tri[:, :, 2] = np.float32(0.25)
    elif kind == "negative":                    # every coordinate below zero: all Morton cells clamp to 0, FLT_MIN max bounds (Q4)
        tri -= np.float32(5.0)
    elif kind == "duplicates":                  # equal keys: one chain per group of 8 identical triangles (Q8)
        tri = np.repeat(tri[: n // 8], 8, axis=0)
    elif kind == "degenerate":                  # zero-area triangles: points and segments among ordinary ones
        tri[::5, 1] = tri[::5, 0]; tri[::5, 2] = tri[::5, 0]
        tri[1::7, 2] = tri[1::7, 1]
    elif kind == "line_x":                      # centroids on a line: two axes collapse
        c = tri.mean(axis=1, keepdims=True)
        tri = tri - c
        c[:, :, 1] = np.float32(0.5); c[:, :, 2] = np.float32(0.5)
        tri = (tri * np.float32(0.6) + c).astype(np.float32)
    elif kind == "mixed_sign":                  # box straddles the origin on every axis
        tri -= np.float32(0.5)
    elif kind == "huge_and_tiny":               # half of the triangles in a 1e-4 cluster of an 8-unit scene: their keys collide
        tri[: n // 2] = tri[: n // 2] * np.float32(1e-4)       # (a larger scene cannot be looked at: the reference adds the camera
        tri[n // 2:] = tri[n // 2:] * np.float32(8.0)          #  position to the ray direction)
    else:
        raise ValueError(kind)
    return np.ascontiguousarray(tri, dtype=np.float32)


def edge_scene(kind, n=600, seed=11, tris_per_mesh=256):
    tri = edge_triangles(kind, n, seed)
    model = []
    for first in range(0, len(tri), tris_per_mesh):
        verts = tri[first:first + tris_per_mesh].reshape(-1, 3)
        model.append((verts, np.arange(len(verts), dtype=np.uint32), 0))
    lights = np.array([[0.5, 10.0, -3.3, 1000.0]], dtype=np.float32)
    return scenes.pack_scene(lights, [model])


def edge_camera(kind, resX, resY, n=600, seed=11):
    """Identity orientation, looking along +z from in front of the geometry's box (the Demo's rule needs a non-zero z extent)."""
    from openclraytracer_b200 import api
    tri = edge_triangles(kind, n, seed).reshape(-1, 3).astype(np.float64)
    lo, hi = tri.min(axis=0), tri.max(axis=0)
    c = (lo + hi) * 0.5
    ext = max((hi - lo).max(), 1e-3)
    pos = np.array([c[0], c[1], lo[2] - 0.75 * ext], dtype=np.float32)
    fov = np.float32(0.5 * resX)                # 90 degrees horizontally
    m = np.array([1, 0, 0, pos[0], 0, 1, 0, pos[1], 0, 0, 1, pos[2]], dtype=np.float32)
    return api.make_camera(resX, resY, fov, m)
