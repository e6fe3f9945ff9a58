"""ctypes face of the CHECKERS: oracle/libclrt_oracle.so (our CPU restatement) and oracle/_ref/libclrt_ref.so (the
reference's own headers compiled in place).  TEST INFRASTRUCTURE — imported only by tests/, __graft_entry__.smoke()
and bench.py's cpu_baseline / --impl reference legs; never by the product package.
"""
import ctypes as C
import os
import subprocess
import sys
import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(_HERE))
from openclraytracer_b200 import layouts as L  # noqa: E402  (layouts only; no CUDA library is loaded)

ORACLE_SO = os.path.join(_HERE, "libclrt_oracle.so")
REF_SO = os.path.join(_HERE, "_ref", "libclrt_ref.so")
REFERENCE_ROOT = "/root/reference/OpenCLRayTracer"

_oracle = None
_ref = None


def build(which=("oracle", "ref")):
    for target in which:
        subprocess.run(["make", "-s", "-C", _HERE, target], check=True)


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def oracle():
    global _oracle
    if _oracle is None:
        if not os.path.exists(ORACLE_SO):
            build(("oracle",))
        _oracle = C.CDLL(ORACLE_SO)
        _oracle.oracle_sort_pairs_bitonic.restype = C.c_int
        _oracle.oracle_num_threads.restype = C.c_int
    return _oracle


def have_ref():
    return os.path.exists(REF_SO)


def ref():
    global _ref
    if _ref is None:
        if not have_ref():
            raise RuntimeError("oracle/_ref/libclrt_ref.so is absent (it can only be built where /root/reference is mounted)")
        _ref = C.CDLL(REF_SO)
        _ref.ref_scene_pack.restype = C.c_int
        _ref.ref_num_threads.restype = C.c_int
    return _ref


# ---- oracle (restatement) ---------------------------------------------------------------------------------------
def flatten(scene):
    n = int(np.frombuffer(scene[:96].tobytes(), dtype=L.SCENE_HEADER)[0]["totalNumberOfTriangles"])
    tris = np.zeros(n, dtype=L.TRIANGLE)
    tim = np.zeros(n, dtype=np.uint32)
    oracle().oracle_flatten(_p(scene), _p(tris), _p(tim))
    return tris, tim


def bvh_morton(scene, tris, tim):
    n = len(tris)
    nodes = np.zeros(max(2 * n - 1, 1), dtype=L.BVH_NODE)
    pairs = np.zeros(n, dtype=L.UINT2)
    oracle().oracle_bvh_morton(_p(scene), _p(tris), _p(tim), C.c_uint32(n), _p(nodes), _p(pairs))
    return nodes, pairs


def sort_pairs_stable(pairs):
    out = np.ascontiguousarray(pairs.copy())
    oracle().oracle_sort_pairs_stable(_p(out), C.c_uint64(len(out)))
    return out


def sort_pairs_bitonic(pairs):
    out = np.ascontiguousarray(pairs.copy())
    rc = oracle().oracle_sort_pairs_bitonic(_p(out), C.c_uint64(len(out)))
    if rc:
        raise ValueError("bitonic sort needs a power-of-two count")
    return out


def scan_inclusive(a):
    a = np.ascontiguousarray(a, dtype=np.uint32)
    out = np.zeros_like(a)
    oracle().oracle_scan_inclusive(_p(a), _p(out), C.c_uint64(len(a)))
    return out


def bvh_build(scene, literal=False):
    """Full oracle build: returns dict(tris, tim, nodes, pairs) — canonical (stable) sort, FIXED node count unless literal."""
    tris, tim = flatten(scene)
    nodes, pairs = bvh_morton(scene, tris, tim)
    pairs = sort_pairs_stable(pairs)
    n = len(tris)
    if n > 1:
        oracle().oracle_bvh_radix_tree(_p(nodes), _p(pairs), C.c_uint32(n), C.c_int(int(literal)))
        oracle().oracle_bvh_refit(_p(nodes), C.c_uint32(n))
    return dict(tris=tris, tim=tim, nodes=nodes, pairs=pairs, n=n)


def bvh_trace_primary(camera, b, want_stats=False):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    count = int(cam["resX"].item()) * int(cam["resY"].item())
    out = np.zeros(count, dtype=L.CONTACT)
    stats = np.zeros(3, dtype=np.uint64)
    oracle().oracle_bvh_trace_primary(_p(cam), _p(b["nodes"]), C.c_uint32(b["n"]), _p(b["tris"]), _p(out), _p(stats) if want_stats else None)
    return (out, stats) if want_stats else out


def bvh_trace_rays(rays, b):
    rays = np.ascontiguousarray(rays, dtype=L.RAY)
    out = np.zeros(len(rays), dtype=L.CONTACT)
    oracle().oracle_bvh_trace_rays(_p(rays), C.c_uint32(len(rays)), _p(b["nodes"]), C.c_uint32(b["n"]), _p(b["tris"]), _p(out))
    return out


def generate_primary_rays(camera):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    rays = np.zeros(int(cam["resX"].item()) * int(cam["resY"].item()), dtype=L.RAY)
    oracle().oracle_generate_primary_rays(_p(cam), _p(rays))
    return rays


def demo_camera(scene, w, h):
    cam = np.zeros((), dtype=L.CAMERA)
    oracle().oracle_demo_camera(_p(scene), C.c_uint32(w), C.c_uint32(h), _p(cam))
    return cam


def camera_from_pose(w, h, fov_deg, quat, pos):
    cam = np.zeros((), dtype=L.CAMERA)
    q = np.asarray(quat, dtype=np.float32); p = np.asarray(pos, dtype=np.float32)
    oracle().oracle_camera_from_pose(C.c_uint32(w), C.c_uint32(h), C.c_float(fov_deg), _p(q), _p(p), _p(cam))
    return cam


def shadow_rays(camera, scene, contacts):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    nl = int(np.frombuffer(scene[:96].tobytes(), dtype=L.SCENE_HEADER)[0]["numberOfLights"])
    rays = np.zeros(len(contacts) * (nl + 1), dtype=L.RAY)
    oracle().oracle_shadow_rays(_p(cam), _p(scene), _p(contacts), C.c_uint32(len(contacts)), _p(rays))
    return rays


def shade(camera, scene, contacts, shadow=None, want_f32=False):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    rgba8 = np.zeros(len(contacts), dtype=np.uint32)
    f32 = np.zeros((len(contacts), 4), dtype=np.float32)
    oracle().oracle_shade(_p(cam), _p(scene), _p(contacts), C.c_uint32(len(contacts)), _p(shadow) if shadow is not None else None,
                          _p(rgba8), _p(f32))
    return (rgba8, f32) if want_f32 else rgba8


# ---- the reference's own code (oracle/_ref) ---------------------------------------------------------------------
def ref_struct_sizes():
    out = np.zeros(14, dtype=np.uint32)
    ref().ref_struct_sizes(_p(out))
    return out


def ref_scene_pack(lights, obj_paths):
    lights = np.ascontiguousarray(lights, dtype=np.float32).reshape(-1, 4)
    arr = (C.c_char_p * len(obj_paths))(*[p.encode() for p in obj_paths])
    buf = C.c_void_p(); size = C.c_uint64()
    rc = ref().ref_scene_pack(_p(lights), C.c_uint32(len(lights)), arr, C.c_uint32(len(obj_paths)), C.byref(buf), C.byref(size))
    if rc:
        raise RuntimeError("reference tinyobj failed to load " + ", ".join(obj_paths))
    out = np.ctypeslib.as_array(C.cast(buf, C.POINTER(C.c_uint8)), shape=(size.value,)).copy()
    ref().ref_free(buf)
    return out


def ref_bvh_build(scene, literal=False, sorter=None):
    """The reference's kernels in launch order; the sort is the canonical stable one unless `sorter` is given."""
    n = int(np.frombuffer(scene[:96].tobytes(), dtype=L.SCENE_HEADER)[0]["totalNumberOfTriangles"])
    nodes = np.zeros(max(2 * n - 1, 1), dtype=L.BVH_NODE)
    pairs = np.zeros(n, dtype=L.UINT2)
    ref().ref_bvh_morton(_p(scene), _p(nodes), _p(pairs))
    unsorted = pairs.copy()
    pairs = (sorter or sort_pairs_stable)(pairs)
    if n > 1:
        ref().ref_bvh_radix_tree(_p(nodes), _p(pairs), C.c_uint32(n), C.c_int(int(literal)))
        ref().ref_bvh_refit(_p(nodes), C.c_uint32(n))
    return dict(nodes=nodes, pairs=pairs, unsorted=unsorted, n=n)


def ref_trace_primary(camera, nodes, n, scene):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    out = np.zeros(int(cam["resX"].item()) * int(cam["resY"].item()), dtype=L.CONTACT)
    ref().ref_bvh_trace_primary(_p(cam), _p(nodes), C.c_uint32(n), _p(scene), _p(out))
    return out


def ref_trace_rays(rays, nodes, n, scene):
    rays = np.ascontiguousarray(rays, dtype=L.RAY)
    out = np.zeros(len(rays), dtype=L.CONTACT)
    ref().ref_bvh_trace_rays(_p(rays), C.c_uint32(len(rays)), _p(nodes), C.c_uint32(n), _p(scene), _p(out))
    return out


def ref_generate_rays(camera):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    rays = np.zeros(int(cam["resX"].item()) * int(cam["resY"].item()), dtype=L.RAY)
    ref().ref_generate_rays(_p(cam), _p(rays))
    return rays


def ref_demo_camera(scene, w, h):
    cam = np.zeros((), dtype=L.CAMERA)
    ref().ref_demo_camera(_p(scene), C.c_uint32(w), C.c_uint32(h), _p(cam))
    return cam


def ref_camera_from_pose(w, h, fov_deg, quat, pos):
    cam = np.zeros((), dtype=L.CAMERA)
    q = np.asarray(quat, dtype=np.float32); p = np.asarray(pos, dtype=np.float32)
    ref().ref_camera_from_pose(C.c_uint32(w), C.c_uint32(h), C.c_float(fov_deg), _p(q), _p(p), _p(cam))
    return cam


def ref_shade(camera, contacts, scene):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    out = np.zeros((len(contacts), 4), dtype=np.float32)
    ref().ref_shade(_p(cam), _p(contacts), C.c_uint32(len(contacts)), _p(scene), _p(out))
    return out


# ---- Two-Level Grid through the reference's own functions (oracle/_ref) -------------------------------------------
class RefTlg:
    """ref_tlg_build: the reference's TwoLevelGrid.h functions in the manager's launch order (FIXED semantics of
    SURVEY.md Q17/Q18/Q21 unless literal)."""

    def __init__(self, scene, topLevelDensity=2.0, leafDensity=2.0, literal=False):
        r = ref()
        r.ref_tlg_build.restype = C.c_void_p
        self.scene = np.ascontiguousarray(scene)
        self.h = C.c_void_p(r.ref_tlg_build(_p(self.scene), C.c_float(topLevelDensity), C.c_float(leafDensity), C.c_int(int(literal))))

    def sizes(self):
        out = np.zeros(4, dtype=np.uint32)
        ref().ref_tlg_sizes(self.h, _p(out))
        return tuple(int(x) for x in out)

    def export(self):
        tc, tp, lc, lp = self.sizes()
        grid = np.zeros(1, dtype=L.GRID_DATA)
        cells = np.zeros(tc, dtype=L.TOP_CELL); tr = np.zeros(tc, dtype=L.UINT2); tpairs = np.zeros(tp, dtype=L.UINT2)
        lr = np.zeros(lc, dtype=L.UINT2); lpairs = np.zeros(lp, dtype=L.UINT2)
        ref().ref_tlg_export(self.h, _p(grid), _p(cells), _p(tr), _p(tpairs), _p(lr), _p(lpairs))
        return dict(grid=grid[0], top_cells=cells, top_ranges=tr, top_pairs=tpairs, leaf_ranges=lr, leaf_pairs=lpairs)

    def trace_primary(self, camera):
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        out = np.zeros(int(cam["resX"].item()) * int(cam["resY"].item()), dtype=L.CONTACT)
        ref().ref_tlg_trace_primary(self.h, _p(cam), _p(self.scene), _p(out))
        return out

    def trace_rays(self, rays):
        rays = np.ascontiguousarray(rays, dtype=L.RAY)
        out = np.zeros(len(rays), dtype=L.CONTACT)
        ref().ref_tlg_trace_rays(self.h, _p(rays), C.c_uint32(len(rays)), _p(self.scene), _p(out))
        return out

    def free(self):
        if self.h:
            ref().ref_tlg_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass
