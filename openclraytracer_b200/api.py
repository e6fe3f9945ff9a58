"""Host-side mirror of the reference's plugin interface over the C ABI (names and argument meaning follow the reference).

  CLExecutionContext  R/Include/OpenCLUtils/CLExecutionContext.h        -> Context
  CLBuffer            R/Include/OpenCLUtils/CLBuffer.h:39-112           -> Buffer
  Scene               R/Include/Scene/Scene.h:66-106                    -> Scene
  BitonicSort         R/Include/Algorithms/Sorting.h:54-85              -> BitonicSort
  PrefixSum           R/Include/Algorithms/PrefixSum.h:48-83            -> PrefixSum
  AccelerationStructureManager / BVHManager / TwoLevelGridManager
                      R/Include/Algorithms/{AccelerationStructureManager.h:52-78,BVHManager.h:59-126,TwoLevelGridManager.h:61-178}

Errors: the reference returns Result + fills Errata&; here a failed call raises ClrtError carrying the Errata message.
The C++ shim with the literal signatures lives in include/clrt/; this module is the Python face used by tests and bench.
"""
import ctypes as C
import weakref
import numpy as np
from . import _lib
from . import layouts as L
from ._lib import ClrtError, check

SORT_RADIX, SORT_BITONIC_COMPAT = 0, 1
TRACE_FAST, TRACE_REFERENCE_ORDER = 0, 1
STAGE_MORTON, STAGE_SORT, STAGE_EMIT, STAGE_REFIT, STAGE_TOTAL, STAGE_TRACE_TREE = 0, 1, 2, 3, 4, 5
TRACE_TREE_AUTO, TRACE_TREE_ALWAYS, TRACE_TREE_NEVER = 0, 1, 2


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p)


class Context:
    """One device + one in-order stream (the reference's context + queue)."""

    def __init__(self, device=0, stream=None):
        self._lib = _lib.load()
        h = C.c_void_p()
        check(self._lib.clrt_ctx_create(int(device), C.c_void_p(stream) if stream else None, C.byref(h)))
        self.handle = h
        self.device = device
        self._children = weakref.WeakSet()      # buffers / scenes / managers created on this context

    def _adopt(self, child):
        self._children.add(child)

    def finish(self):                      # CLExecutionContext::finishQueue
        check(self._lib.clrt_ctx_sync(self.handle))

    flushQueue = finish

    @property
    def stream(self):
        return self._lib.clrt_ctx_stream(self.handle)

    def enqueueWriteBuffer(self, dptr, arr):           # blocking, like the reference's
        arr = np.ascontiguousarray(arr)
        check(self._lib.clrt_memcpy_htod(self.handle, C.c_void_p(dptr), _ptr(arr), arr.nbytes))

    def enqueueReadBuffer(self, dptr, out):
        assert out.flags["C_CONTIGUOUS"]
        check(self._lib.clrt_memcpy_dtoh(self.handle, _ptr(out), C.c_void_p(dptr), out.nbytes))
        return out

    def device_info(self, what):
        v = C.c_uint64()
        check(self._lib.clrt_ctx_device_info(self.handle, what, C.byref(v)))
        return v.value

    def getMaxComputeUnits(self):
        return self.device_info(2)

    def getMaxWorkGroupSize(self):
        return self.device_info(1)

    def getLocalMemSize(self):
        return self.device_info(0)

    def close(self):
        """Frees everything still alive on this context first (handles keep a raw pointer to their context)."""
        if self.handle:
            kids = list(self._children)
            for k in sorted(kids, key=lambda k: 0 if isinstance(k, AccelerationStructureManager) else 1):
                k.free()
            self._lib.clrt_ctx_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class Buffer:
    """CLBuffer: device allocation with grow-only resize and blocking copyToHost."""

    def __init__(self, context, size, source=None):
        self._lib = _lib.load()
        self.context = context
        h = C.c_void_p()
        src = None
        if source is not None:
            source = np.ascontiguousarray(source)
            assert source.nbytes >= size
            src = _ptr(source)
        check(self._lib.clrt_buf_alloc(context.handle, size, src, C.byref(h)))
        self.handle = h
        context._adopt(self)

    @classmethod
    def from_array(cls, context, arr):
        arr = np.ascontiguousarray(arr)
        return cls(context, arr.nbytes, arr)

    def resize(self, new_size):
        check(self._lib.clrt_buf_resize(self.handle, new_size))

    def getSize(self):
        return self._lib.clrt_buf_size(self.handle)

    def getActualSize(self):
        return self._lib.clrt_buf_capacity(self.handle)

    def getCLMem(self):
        return self._lib.clrt_buf_dptr(self.handle)

    dptr = property(getCLMem)

    def fill(self, pattern, nbytes=None):              # CLExecutionContext::enqueueFillBuffer
        pattern = np.ascontiguousarray(pattern)
        check(self._lib.clrt_buf_fill(self.handle, _ptr(pattern), pattern.nbytes, self.getSize() if nbytes is None else nbytes))

    def upload(self, arr, offset=0):
        arr = np.ascontiguousarray(arr)
        check(self._lib.clrt_buf_upload(self.handle, _ptr(arr), offset, arr.nbytes))

    def copyToHost(self, dtype=np.uint8, count=None, offset=0):
        dtype = np.dtype(dtype)
        n = (self.getSize() - offset) // dtype.itemsize if count is None else count
        out = np.empty(n, dtype=dtype)
        check(self._lib.clrt_buf_download(self.handle, _ptr(out), offset, out.nbytes))
        return out

    def free(self):
        if self.handle:
            self._lib.clrt_buf_free(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class BitonicSort:
    """Common::BitonicSort(context, useKeyValue): sort(cl_mem, n) sorts n (key,value) pairs in place by key."""

    def __init__(self, context, useKeyValue=True, mode=SORT_RADIX):
        self._lib = _lib.load()
        self.context = context
        self.mode = mode

    def initialize(self):
        return 0                                           # kernels are compiled ahead of time

    def sort(self, cl_mem, num_items):
        check(self._lib.clrt_sort_pairs_u32(self.context.handle, C.c_void_p(cl_mem), num_items, self.mode))


class PrefixSum:
    """Common::PrefixSum(context): computePrefixSum(in, out, n) — inclusive scan of uint32."""

    def __init__(self, context):
        self._lib = _lib.load()
        self.context = context

    def initialize(self):
        return 0

    def computePrefixSum(self, in_mem, out_mem, n):
        check(self._lib.clrt_scan_inclusive_u32(self.context.handle, C.c_void_p(in_mem), C.c_void_p(out_mem), n))


class Scene:
    """Scene: host buffer in the packed v0.6 format + loadToGPU."""

    def __init__(self, host_scene_data=None):
        self._lib = _lib.load()
        self._host = None if host_scene_data is None else np.ascontiguousarray(host_scene_data, dtype=np.uint8)
        self.handle = None
        self.context = None

    def getHostSceneData(self):
        return self._host

    def getSceneDataSize(self):
        return 0 if self._host is None else self._host.nbytes

    def loadToGPU(self, context):
        h = C.c_void_p()
        check(self._lib.clrt_scene_upload(context.handle, _ptr(self._host), self._host.nbytes, C.byref(h)))
        self.handle, self.context = h, context
        context._adopt(self)

    def getDeviceSceneData(self):
        return self._lib.clrt_scene_dptr(self.handle)

    def triangle_count(self):
        return int(self._lib.clrt_scene_triangle_count(self.handle))

    def export_triangles(self):
        out = np.empty(self.triangle_count(), dtype=L.TRIANGLE)
        check(self._lib.clrt_scene_export_triangles(self.handle, _ptr(out)))
        return out

    def header(self):
        return np.frombuffer(self._host[:96].tobytes(), dtype=L.SCENE_HEADER)[0]

    def num_lights(self):
        return int(self.header()["numberOfLights"])

    def free(self):
        if self.handle:
            self._lib.clrt_scene_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


def make_camera(resX, resY, FOVDistance, viewTransform):
    cam = np.zeros((), dtype=L.CAMERA)
    cam["FOVDistance"] = FOVDistance
    cam["resX"], cam["resY"], cam["supersampilngFactor"] = resX, resY, 1
    cam["viewTransform"] = np.asarray(viewTransform, dtype=np.float32).reshape(12)
    return cam


def demo_camera(scene_buf, resX, resY, fov_deg=90.0):
    """The Demo's camera rule (OpenCLRayTracerDemo.cpp:202-217): FOV 90 deg, identity orientation,
    position (cx, cy, minZ - halfZ) of the scene box.  float32 arithmetic as the reference."""
    hdr = np.frombuffer(scene_buf[:96].tobytes(), dtype=L.SCENE_HEADER)[0]
    bb = hdr["modelsBoundingBox"]
    f = np.float32
    half = (bb[1, :3] - bb[0, :3]) * f(0.5)
    pos = np.array([bb[0, 0] + half[0], bb[0, 1] + half[1], bb[0, 2] - half[2]], dtype=np.float32)
    deg2rad = f(f(3.14159265359) / f(180.0))
    fovd = f(f(resY) * f(0.5)) / f(np.tan(f(f(fov_deg) * f(0.5) * deg2rad), dtype=np.float32))
    m = np.array([1, 0, 0, pos[0], 0, 1, 0, pos[1], 0, 0, 1, pos[2]], dtype=np.float32)
    return make_camera(resX, resY, f(fovd), m)


class AccelerationStructureManager:
    """Abstract base: initialize / initializeFrame / construct / generateContacts(camera) /
    generateContacts(rays, contacts, rayCount) / getPrimaryContacts."""

    def __init__(self, context, scene):
        self._lib = _lib.load()
        self.context, self.scene = context, scene
        self.handle = None
        self._primary_pixels = 0

    def initialize(self):
        raise NotImplementedError

    def initializeFrame(self):
        check(self._lib.clrt_accel_init_frame(self.handle))

    def construct(self):
        check(self._lib.clrt_accel_construct(self.handle))

    def generateContacts(self, camera_or_rays, contacts=None, rayCount=None):
        if rayCount is None:                                  # generateContacts(Camera&)
            cam = np.ascontiguousarray(camera_or_rays, dtype=L.CAMERA)
            out = None if contacts is None else C.c_void_p(contacts.getCLMem() if isinstance(contacts, Buffer) else contacts)
            check(self._lib.clrt_trace_primary(self.handle, _ptr(cam), out))
            self._primary_pixels = int(cam["resX"].item()) * int(cam["resY"].item())
        else:                                                 # generateContacts(rays, contacts, rayCount)
            r = camera_or_rays.getCLMem() if isinstance(camera_or_rays, Buffer) else camera_or_rays
            c = contacts.getCLMem() if isinstance(contacts, Buffer) else contacts
            check(self._lib.clrt_trace_rays(self.handle, C.c_void_p(r), C.c_void_p(c), rayCount))

    def trace_primary_stripe(self, camera, rank, nranks, dptr_contacts):
        """generateContacts(Camera&) for one image-space stripe, contacts in stripe order (clrt_trace_primary_stripe)."""
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        check(self._lib.clrt_trace_primary_stripe(self.handle, _ptr(cam), rank, nranks, C.c_void_p(dptr_contacts)))

    def getPrimaryContacts(self):
        """Device pointer of the primary contacts (the reference returns the CLBuffer)."""
        return self._lib.clrt_accel_primary_contacts_dptr(self.handle)

    def check(self):
        """Waits for the context's stream; raises if a traversal stack overflowed since the last check (clrt_accel_check)."""
        check(self._lib.clrt_accel_check(self.handle))

    def download_primary_contacts(self):
        """getPrimaryContacts()->copyToHost (OpenCLRayTracerDemo.cpp:289)."""
        self.check()
        out = np.empty(self._primary_pixels, dtype=L.CONTACT)
        self.context.enqueueReadBuffer(self.getPrimaryContacts(), out)
        return out

    def set_trace_mode(self, mode):
        check(self._lib.clrt_accel_set_trace_mode(self.handle, mode))

    def set_trace_tree(self, policy):                   # 0 auto, 1 always, 2 never; effective at the next construct()
        check(self._lib.clrt_accel_set_trace_tree(self.handle, policy))

    def has_trace_tree(self):
        import ctypes
        v = ctypes.c_int(0)
        check(self._lib.clrt_accel_has_trace_tree(self.handle, ctypes.byref(v)))
        return bool(v.value)

    def enable_stage_timing(self, on=True):
        check(self._lib.clrt_accel_enable_stage_timing(self.handle, int(on)))

    def stage_ms(self, stage):
        v = C.c_float()
        check(self._lib.clrt_accel_stage_ms(self.handle, stage, C.byref(v)))
        return v.value

    def shadow_shade(self, camera, dptr_contacts, n, dptr_rgba8):
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        check(self._lib.clrt_shadow_shade(self.handle, _ptr(cam), C.c_void_p(dptr_contacts), n, C.c_void_p(dptr_rgba8)))

    def set_fused_shadows(self, on=True):
        check(self._lib.clrt_accel_set_fused_shadows(self.handle, int(on)))

    def set_stats(self, on=True):
        check(self._lib.clrt_accel_set_stats(self.handle, int(on)))

    def read_stats(self):
        """{rays traversed, inner-node visits, triangle tests, run leaf-box tests} since set_stats(True)."""
        v = (C.c_uint64 * 4)()
        check(self._lib.clrt_accel_read_stats(self.handle, v))
        return int(v[0]), int(v[1]), int(v[2]), int(v[3])

    def render_stripe(self, camera, shadows, rank, nranks, dptr_rgba8, want_rays=False):
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        rays = C.c_uint64()
        check(self._lib.clrt_render_stripe(self.handle, _ptr(cam), int(shadows), rank, nranks, C.c_void_p(dptr_rgba8),
                                           C.byref(rays) if want_rays else None))
        return rays.value

    def render_stripe_on(self, stream, slot, camera, shadows, rank, nranks, dptr_rgba8, want_rays=False):
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        rays = C.c_uint64()
        check(self._lib.clrt_render_stripe_on(self.handle, C.c_void_p(stream), slot, _ptr(cam), int(shadows), rank, nranks,
                                              C.c_void_p(dptr_rgba8), C.byref(rays) if want_rays else None))
        return rays.value

    def render_stripe_peers(self, stream, slot, camera, rank, nranks, peer_rasters, want_rays=False):
        """peer_rasters: ctypes array of nranks device pointers (peer_pointer_array)."""
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        rays = C.c_uint64()
        check(self._lib.clrt_render_stripe_peers(self.handle, C.c_void_p(stream), slot, _ptr(cam), rank, nranks, peer_rasters,
                                                 C.byref(rays) if want_rays else None))
        return rays.value

    def render_frame_peers(self, stream, slot, camera, rank, nranks, peer_rasters, peer_words, seq, h_rgba8=None, want_rays=False,
                           consumer_stream=None):
        """The whole per-frame protocol of the fused peer-memory exchange in one call (include/clrt.h)."""
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        rays = C.c_uint64()
        check(self._lib.clrt_render_frame_peers(self.handle, C.c_void_p(stream), slot, _ptr(cam), rank, nranks, peer_rasters, peer_words, seq,
                                                C.c_void_p(h_rgba8) if h_rgba8 else None, C.byref(rays) if want_rays else None,
                                                C.c_void_p(consumer_stream) if consumer_stream else None))
        return rays.value

    def render_bounce_stripe_on(self, stream, slot, camera, rank, nranks, seed, dptr_contacts, dptr_rays, dptr_bounce, dptr_rgba8):
        """configs[3] frame of one rank on a frame slot: primary stripe -> diffuse bounce rays -> generic-ray walk -> shading (clrt_render_bounce_stripe_on)."""
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        check(self._lib.clrt_render_bounce_stripe_on(self.handle, C.c_void_p(stream), slot, _ptr(cam), rank, nranks, seed, C.c_void_p(dptr_contacts),
                                                     C.c_void_p(dptr_rays), C.c_void_p(dptr_bounce), C.c_void_p(dptr_rgba8)))

    def render_stripes_batch(self, stream, slot, cameras, rank, nranks, dptr_rgba8):
        """1..4 stripes (one camera each) by one pair of launches into a local stripe-ordered buffer (clrt_render_stripes_batch)."""
        cams = np.ascontiguousarray(np.stack([np.asarray(c, dtype=L.CAMERA).reshape(()) for c in cameras]))
        check(self._lib.clrt_render_stripes_batch(self.handle, C.c_void_p(stream), slot, len(cameras), _ptr(cams), rank, nranks, C.c_void_p(dptr_rgba8)))

    def render_frames_peers(self, stream, slot0, cameras, rank, nranks, peer_rasters_per_frame, peer_words, seqs, h_rgba8s=None, consumer_stream=None):
        """1..4 consecutive frames of the fused exchange traced by one pair of launches (clrt_render_frames_peers).
        peer_rasters_per_frame[f] = the pointer array of frame f's framebuffers on every GPU (peer_pointer_array)."""
        k = len(cameras)
        cams = np.ascontiguousarray(np.stack([np.asarray(c, dtype=L.CAMERA).reshape(()) for c in cameras]))
        rasters = (C.c_void_p * (k * 8))()
        for f in range(k):
            for g in range(nranks):
                rasters[f * 8 + g] = peer_rasters_per_frame[f][g]
        sq = (C.c_uint32 * k)(*seqs)
        hosts = None
        if h_rgba8s is not None:
            hosts = (C.c_void_p * k)(*[C.c_void_p(h) if h else None for h in h_rgba8s])
        check(self._lib.clrt_render_frames_peers(self.handle, C.c_void_p(stream), slot0, k, _ptr(cams), rank, nranks, rasters, peer_words, sq, hosts,
                                                 C.c_void_p(consumer_stream)))

    def render_frame_host(self, camera, shadows, out_rgba8=None, want_rays=False):
        cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
        if out_rgba8 is None:
            out_rgba8 = np.empty((int(cam["resY"].item()), int(cam["resX"].item()), 4), dtype=np.uint8)
        rays = C.c_uint64()
        ptr = out_rgba8 if isinstance(out_rgba8, int) else _ptr(out_rgba8)
        check(self._lib.clrt_render_frame_host(self.handle, _ptr(cam), int(shadows), C.c_void_p(ptr) if isinstance(ptr, int) else ptr,
                                               C.byref(rays) if want_rays else None))
        return out_rgba8, rays.value

    def free(self):
        if self.handle:
            self._lib.clrt_accel_destroy(self.handle)
            self.handle = None

    def __del__(self):
        try:
            self.free()
        except Exception:
            pass


class BVHManager(AccelerationStructureManager):
    def initialize(self):
        h = C.c_void_p()
        check(self._lib.clrt_bvh_create(self.context.handle, self.scene.handle, C.byref(h)))
        self.handle = h
        self.context._adopt(self)

    def export_nodes(self):
        n = self.scene.triangle_count()
        out = np.zeros(max(2 * n - 1, 0), dtype=L.BVH_NODE)
        check(self._lib.clrt_bvh_export_nodes(self.handle, _ptr(out)))
        return out

    def export_pairs(self):
        out = np.zeros(self.scene.triangle_count(), dtype=L.UINT2)
        check(self._lib.clrt_bvh_export_pairs(self.handle, _ptr(out)))
        return out


class TwoLevelGridManager(AccelerationStructureManager):
    def __init__(self, context, scene, topLevelDensity=2.0, leafDensity=2.0):
        super().__init__(context, scene)
        self._top, self._leaf = float(topLevelDensity), float(leafDensity)

    def setTopLevelDensity(self, v):                 # read at the next initializeFrame(), like the reference's calculateGridData
        self._top = float(v)
        if self.handle:
            check(self._lib.clrt_tlg_set_densities(self.handle, self._top, self._leaf))

    def setLeafDensity(self, v):
        self._leaf = float(v)
        if self.handle:
            check(self._lib.clrt_tlg_set_densities(self.handle, self._top, self._leaf))

    def getResolution(self):                         # TwoLevelGridManager.h:129
        v = (C.c_uint32 * 3)()
        check(self._lib.clrt_tlg_resolution(self.handle, v))
        return tuple(v)

    def getTopLevelDensity(self):
        return self._top

    def getLeafDensity(self):
        return self._leaf

    def initialize(self):
        h = C.c_void_p()
        check(self._lib.clrt_tlg_create(self.context.handle, self.scene.handle, self._top, self._leaf, C.byref(h)))
        self.handle = h
        self.context._adopt(self)

    def sizes(self):
        v = [C.c_uint32() for _ in range(4)]
        check(self._lib.clrt_tlg_sizes(self.handle, *[C.byref(x) for x in v]))
        return tuple(x.value for x in v)

    def export(self):
        top_cells, top_pairs, leaf_cells, leaf_pairs = self.sizes()
        grid = np.zeros((), dtype=L.GRID_DATA)
        cells = np.zeros(top_cells, dtype=L.TOP_CELL)
        top_ranges = np.zeros(top_cells, dtype=L.UINT2)
        tpairs = np.zeros(top_pairs, dtype=L.UINT2)
        leaf_ranges = np.zeros(leaf_cells, dtype=L.UINT2)
        lpairs = np.zeros(leaf_pairs, dtype=L.UINT2)
        check(self._lib.clrt_tlg_export(self.handle, _ptr(grid), _ptr(cells), _ptr(top_ranges), _ptr(tpairs), _ptr(leaf_ranges), _ptr(lpairs)))
        return dict(grid=grid, top_cells=cells, top_ranges=top_ranges, top_pairs=tpairs, leaf_ranges=leaf_ranges, leaf_pairs=lpairs)


class KDTreeManager(AccelerationStructureManager):
    """The reference's TODO 2.1 (KNOWN_ISSUES_AND_TODO.txt:24-26) as a third AccelerationStructureManager (csrc/kd.cu)."""

    def initialize(self):
        h = C.c_void_p()
        check(self._lib.clrt_kdtree_create(self.context.handle, self.scene.handle, C.byref(h)))
        self.handle = h
        self.context._adopt(self)

    def info(self):
        v = (C.c_uint32 * 3)(); f = (C.c_float * 6)()
        check(self._lib.clrt_kd_info(self.handle, v, f))
        return dict(levels=v[0], pairs=v[1], leaves=v[2], origin=tuple(f[:3]), cell=tuple(f[3:]))


def generate_primary_rays(context, camera, dptr_rays):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    check(_lib.load().clrt_generate_primary_rays(context.handle, _ptr(cam), C.c_void_p(dptr_rays)))


def generate_shadow_rays(context, scene, camera, dptr_contacts, n, dptr_rays):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    per = C.c_uint32()
    check(_lib.load().clrt_generate_shadow_rays(context.handle, scene.handle, _ptr(cam), C.c_void_p(dptr_contacts), n,
                                                C.c_void_p(dptr_rays), C.byref(per)))
    return per.value


def generate_bounce_rays(context, camera, dptr_contacts, n, seed, dptr_rays):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    check(_lib.load().clrt_generate_bounce_rays(context.handle, _ptr(cam), C.c_void_p(dptr_contacts), n, C.c_uint64(seed), C.c_void_p(dptr_rays)))


RAYS_DIFFUSE, RAYS_REFLECT = 0, 1


def generate_secondary_rays(context, kind, camera, dptr_in_rays, dptr_contacts, n, seed, dptr_rays):
    """camera: the primary pass's camera (contacts of generateContacts(camera)) or None with dptr_in_rays (contacts of
    generateContacts(rays, contacts, n) over those rays) — include/clrt.h."""
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA) if camera is not None else None
    check(_lib.load().clrt_generate_secondary_rays(context.handle, kind, _ptr(cam) if cam is not None else None,
                                                   C.c_void_p(dptr_in_rays) if dptr_in_rays else None, C.c_void_p(dptr_contacts), n,
                                                   C.c_uint64(seed), C.c_void_p(dptr_rays)))


def shade(context, scene, camera, dptr_contacts, n, dptr_shadow=None, dptr_rgba8=None, dptr_rgba_f32=None):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    check(_lib.load().clrt_shade(context.handle, scene.handle, _ptr(cam), C.c_void_p(dptr_contacts), n,
                                 C.c_void_p(dptr_shadow) if dptr_shadow else None,
                                 C.c_void_p(dptr_rgba8) if dptr_rgba8 else None,
                                 C.c_void_p(dptr_rgba_f32) if dptr_rgba_f32 else None))


def shade_bounce(context, scene, camera, dptr_contacts, n, dptr_bounce_contacts, dptr_rgba8):
    cam = np.ascontiguousarray(camera, dtype=L.CAMERA)
    check(_lib.load().clrt_shade_bounce(context.handle, scene.handle, _ptr(cam), C.c_void_p(dptr_contacts), n,
                                        C.c_void_p(dptr_bounce_contacts), C.c_void_p(dptr_rgba8)))


def frame_untile(context, dptr_stripes, resX, resY, nranks, dptr_raster):
    check(_lib.load().clrt_frame_untile(context.handle, C.c_void_p(dptr_stripes), resX, resY, nranks, C.c_void_p(dptr_raster)))


def frame_untile_on(context, stream, dptr_stripes, resX, resY, nranks, dptr_raster):
    check(_lib.load().clrt_frame_untile_on(context.handle, C.c_void_p(stream), C.c_void_p(dptr_stripes), resX, resY, nranks,
                                           C.c_void_p(dptr_raster)))


# ---- frame exchange over peer memory (include/clrt.h, csrc/peer.cu) -----------------------------------------------------------
PEER_HANDLE_BYTES = 64


def peer_pointer_array(ptrs):
    return (C.c_void_p * len(ptrs))(*[C.c_void_p(int(p)) for p in ptrs])


def peer_alloc(context, nbytes):
    p = C.c_void_p()
    check(_lib.load().clrt_peer_alloc(context.handle, nbytes, C.byref(p)))
    return p.value


def peer_free(context, dptr):
    check(_lib.load().clrt_peer_free(context.handle, C.c_void_p(dptr)))


def peer_export(context, dptr):
    buf = C.create_string_buffer(PEER_HANDLE_BYTES)
    check(_lib.load().clrt_peer_export(context.handle, C.c_void_p(dptr), buf))
    return buf.raw


def peer_open(context, handle):
    assert len(handle) == PEER_HANDLE_BYTES
    p = C.c_void_p()
    check(_lib.load().clrt_peer_open(context.handle, handle, C.byref(p)))
    return p.value


def peer_close(context, dptr):
    check(_lib.load().clrt_peer_close(context.handle, C.c_void_p(dptr)))


def frame_push(context, stream, dptr_stripe, resX, resY, rank, nranks, peer_rasters):
    check(_lib.load().clrt_frame_push(context.handle, C.c_void_p(stream), C.c_void_p(dptr_stripe), resX, resY, rank, nranks, peer_rasters))


def peer_signal(context, stream, peer_words, npeers, word, value):
    check(_lib.load().clrt_peer_signal(context.handle, C.c_void_p(stream), peer_words, npeers, word, value))


def peer_wait(context, stream, dptr_words, first_word, n, value, timeout_ms=0):
    check(_lib.load().clrt_peer_wait(context.handle, C.c_void_p(stream), C.c_void_p(dptr_words), first_word, n, value, timeout_ms))


def frame_stripe_pixels(resX, resY, rank, nranks):
    px, tiles = C.c_uint32(), C.c_uint32()
    check(_lib.load().clrt_frame_stripe_pixels(resX, resY, rank, nranks, C.byref(px), C.byref(tiles)))
    return px.value, tiles.value
