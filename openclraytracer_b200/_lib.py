"""ctypes binding of libclrt_cuda.so (include/clrt.h).

The library is the product; there is NO fallback: if it is missing the import fails loudly and tells how to build it.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("CLRT_LIB_PATH") or os.path.join(_HERE, "libclrt_cuda.so")   # CLRT_LIB_PATH: kernel-tuning builds

c_void_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); must list every symbol declared in include/clrt.h (tests/test_abi.py checks it)
SIGNATURES = {
    "clrt_last_error": (C.c_char_p, []),
    "clrt_abi_version": (C.c_int, []),
    "clrt_ctx_create": (C.c_int, [C.c_int, C.c_void_p, c_void_pp]),
    "clrt_ctx_destroy": (C.c_int, [C.c_void_p]),
    "clrt_ctx_sync": (C.c_int, [C.c_void_p]),
    "clrt_ctx_stream": (C.c_void_p, [C.c_void_p]),
    "clrt_ctx_device_info": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_uint64)]),
    "clrt_buf_alloc": (C.c_int, [C.c_void_p, C.c_size_t, C.c_void_p, c_void_pp]),
    "clrt_buf_free": (C.c_int, [C.c_void_p]),
    "clrt_buf_resize": (C.c_int, [C.c_void_p, C.c_size_t]),
    "clrt_buf_size": (C.c_size_t, [C.c_void_p]),
    "clrt_buf_capacity": (C.c_size_t, [C.c_void_p]),
    "clrt_buf_dptr": (C.c_void_p, [C.c_void_p]),
    "clrt_buf_fill": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]),
    "clrt_buf_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]),
    "clrt_buf_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t]),
    "clrt_memcpy_htod": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "clrt_memcpy_dtoh": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "clrt_sort_pairs_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int]),
    "clrt_scan_inclusive_u32": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "clrt_scan_inclusive_u32_noalloc": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_size_t]),
    "clrt_scene_upload": (C.c_int, [C.c_void_p, C.c_void_p, C.c_size_t, c_void_pp]),
    "clrt_scene_destroy": (C.c_int, [C.c_void_p]),
    "clrt_scene_triangle_count": (C.c_uint64, [C.c_void_p]),
    "clrt_scene_dptr": (C.c_void_p, [C.c_void_p]),
    "clrt_scene_triangles_dptr": (C.c_void_p, [C.c_void_p]),
    "clrt_scene_export_triangles": (C.c_int, [C.c_void_p, C.c_void_p]),
    "clrt_bvh_create": (C.c_int, [C.c_void_p, C.c_void_p, c_void_pp]),
    "clrt_tlg_create": (C.c_int, [C.c_void_p, C.c_void_p, C.c_float, C.c_float, c_void_pp]),
    "clrt_accel_destroy": (C.c_int, [C.c_void_p]),
    "clrt_accel_init_frame": (C.c_int, [C.c_void_p]),
    "clrt_accel_construct": (C.c_int, [C.c_void_p]),
    "clrt_trace_primary": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "clrt_trace_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32]),
    "clrt_accel_primary_contacts_dptr": (C.c_void_p, [C.c_void_p]),
    "clrt_accel_set_trace_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "clrt_accel_set_trace_tree": (C.c_int, [C.c_void_p, C.c_int]),
    "clrt_accel_has_trace_tree": (C.c_int, [C.c_void_p, C.POINTER(C.c_int)]),
    "clrt_accel_stage_ms": (C.c_int, [C.c_void_p, C.c_int, C.POINTER(C.c_float)]),
    "clrt_accel_enable_stage_timing": (C.c_int, [C.c_void_p, C.c_int]),
    "clrt_accel_set_fused_shadows": (C.c_int, [C.c_void_p, C.c_int]),
    "clrt_accel_set_stats": (C.c_int, [C.c_void_p, C.c_int]),
    "clrt_accel_read_stats": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint64)]),
    "clrt_accel_check": (C.c_int, [C.c_void_p]),
    "clrt_kdtree_create": (C.c_int, [C.c_void_p, C.c_void_p, c_void_pp]),
    "clrt_kd_info": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32), C.POINTER(C.c_float)]),
    "clrt_tlg_set_densities": (C.c_int, [C.c_void_p, C.c_float, C.c_float]),
    "clrt_tlg_resolution": (C.c_int, [C.c_void_p, C.POINTER(C.c_uint32)]),
    "clrt_bvh_export_nodes": (C.c_int, [C.c_void_p, C.c_void_p]),
    "clrt_bvh_export_pairs": (C.c_int, [C.c_void_p, C.c_void_p]),
    "clrt_tlg_sizes": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_uint32)] * 4),
    "clrt_tlg_export": (C.c_int, [C.c_void_p] + [C.c_void_p] * 6),
    "clrt_generate_primary_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
    "clrt_generate_shadow_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint32)]),
    "clrt_generate_bounce_rays": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p]),
    "clrt_generate_secondary_rays": (C.c_int, [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_void_p]),
    "clrt_shade_bounce": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
    "clrt_render_bounce_stripe_on": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint64,
                                              C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "clrt_trace_primary_stripe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
    "clrt_shade": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]),
    "clrt_shadow_shade": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
    "clrt_frame_stripe_pixels": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_uint32), C.POINTER(C.c_uint32)]),
    "clrt_render_stripe": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]),
    "clrt_render_stripe_on": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_uint32, C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64)]),
    "clrt_frame_untile": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
    "clrt_frame_untile_on": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
    "clrt_render_frame_host": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.POINTER(C.c_uint64)]),
    "clrt_peer_alloc": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(C.c_void_p)]),
    "clrt_peer_free": (C.c_int, [C.c_void_p, C.c_void_p]),
    "clrt_peer_export": (C.c_int, [C.c_void_p, C.c_void_p, C.c_char_p]),
    "clrt_peer_open": (C.c_int, [C.c_void_p, C.c_char_p, C.POINTER(C.c_void_p)]),
    "clrt_peer_close": (C.c_int, [C.c_void_p, C.c_void_p]),
    "clrt_render_stripe_peers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_uint64)]),
    "clrt_render_frame_peers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p), C.POINTER(C.c_void_p),
                                          C.c_uint32, C.c_void_p, C.POINTER(C.c_uint64), C.c_void_p]),
    "clrt_render_stripes_batch": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.c_void_p]),
    "clrt_render_frames_peers": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p),
                                           C.POINTER(C.c_void_p), C.POINTER(C.c_uint32), C.POINTER(C.c_void_p), C.c_void_p]),
    "clrt_frame_push": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p)]),
    "clrt_peer_signal": (C.c_int, [C.c_void_p, C.c_void_p, C.POINTER(C.c_void_p), C.c_uint32, C.c_uint32, C.c_uint32]),
    "clrt_peer_wait": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32]),
}

_lib = None


def load():
    """Load libclrt_cuda.so (no compute happens here; safe without a GPU)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(or `make -C openclraytracer_b200/csrc`). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError here = ABI drift; let it surface
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class ClrtError(RuntimeError):
    """Result::Error + Errata (R/Include/Common/Errata.h:47-110) surfaced as an exception."""


def check(rc):
    if rc != 0:
        raise ClrtError(load().clrt_last_error().decode("utf-8", "replace"))
