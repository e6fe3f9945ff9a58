/* clrt.h — C ABI of libclrt_cuda.so: the B200-native (sm_100a) replacement for the hot path of
 * kotturtech/OpenCLRayTracer, sitting where the reference's OpenCL wrapper + kernels sit.
 *
 * R = /root/reference/OpenCLRayTracer.  The reference has no FFI today; every entry point below names the
 * C++ interface (file:line) it stands behind.  The C++ shim in include/clrt/ (namespace CLRayTracer) keeps
 * the reference's class names and signatures on top of these calls; INTEGRATION.md shows the binding.
 *
 * Conventions
 *  - plain pointers and sizes only; opaque handles; no C++/torch types.
 *  - every call returns 0 on success, non-zero on error; the message is in clrt_last_error() (thread local).
 *    This is the reference's `Result {Success=0, Error=1}` + `Errata&` convention (R/Include/Common/Errata.h:47-110).
 *  - a context owns (or borrows) ONE CUDA stream: the reference's single in-order queue
 *    (R/src/OpenCLUtils/CLExecutionContext.cpp:264-293).  Calls enqueue on that stream.  Entry points that
 *    move data to/from HOST memory block until the copy is done (the reference is fully synchronous,
 *    e.g. R/src/CLRayTracerEngine/BVHManager.cpp:184-192); device-only entry points are asynchronous on the
 *    stream — call clrt_ctx_sync() (== clFinish) where the reference semantics are wanted.
 *  - "dptr" arguments are DEVICE pointers (what `cl_mem` becomes); "h" arguments are HOST pointers.
 *  - there is no CPU fallback: without a CUDA device every call fails with an error.
 */
#ifndef CLRT_H
#define CLRT_H
#include <stddef.h>
#include <stdint.h>
#include "clrt_types.h"

#ifdef __cplusplus
extern "C" {
#endif

#define CLRT_ABI_VERSION 1

typedef struct clrt_ctx   clrt_ctx;    /* CLExecutionContext  R/Include/OpenCLUtils/CLExecutionContext.h */
typedef struct clrt_buf   clrt_buf;    /* CLBuffer            R/Include/OpenCLUtils/CLBuffer.h:39-112    */
typedef struct clrt_scene clrt_scene;  /* Scene (device side) R/Include/Scene/Scene.h:66-106             */
typedef struct clrt_accel clrt_accel;  /* AccelerationStructureManager R/Include/Algorithms/AccelerationStructureManager.h:52-78 */

/* ---- errors / version ------------------------------------------------------------------------- */
const char* clrt_last_error(void);                    /* Errata::message()  R/Include/Common/Errata.h:84 */
int         clrt_abi_version(void);

/* ---- context: CLExecutionContext (R/src/OpenCLUtils/CLExecutionContext.cpp:264-293) ----------------- */
/* stream == NULL: the context creates and owns a non-blocking stream; otherwise it borrows the given
 * cudaStream_t (e.g. torch's current stream) and never destroys it. */
int clrt_ctx_create(int device, void* stream, clrt_ctx** out);
int clrt_ctx_destroy(clrt_ctx* ctx);
int clrt_ctx_sync(clrt_ctx* ctx);                     /* CLExecutionContext::finishQueue                */
void* clrt_ctx_stream(clrt_ctx* ctx);                 /* the cudaStream_t                               */
/* Device getters the managers query (R/Include/OpenCLUtils/CLInterface.h:183,284,343). */
#define CLRT_DEV_LOCAL_MEM_SIZE      0   /* getLocalMemSize      -> max shared memory per block (opt-in) */
#define CLRT_DEV_MAX_WORK_GROUP_SIZE 1   /* getMaxWorkGroupSize  -> max threads per block               */
#define CLRT_DEV_MAX_COMPUTE_UNITS   2   /* getMaxComputeUnits   -> SM count                            */
#define CLRT_DEV_GLOBAL_MEM_SIZE     3
int clrt_ctx_device_info(clrt_ctx* ctx, int what, uint64_t* out);

/* ---- buffers: CLBuffer (R/Include/OpenCLUtils/CLBuffer.h:39-112, R/src/OpenCLUtils/CLBuffer.cpp:42-122) -- */
int    clrt_buf_alloc(clrt_ctx* ctx, size_t bytes, const void* h_src /* may be NULL */, clrt_buf** out);
int    clrt_buf_free(clrt_buf* buf);
int    clrt_buf_resize(clrt_buf* buf, size_t bytes);  /* grow-only; contents discarded on growth (CLBuffer.cpp:70-94) */
size_t clrt_buf_size(const clrt_buf* buf);            /* getSize        */
size_t clrt_buf_capacity(const clrt_buf* buf);        /* getActualSize  */
void*  clrt_buf_dptr(const clrt_buf* buf);            /* getCLMem       */
/* CLExecutionContext::enqueueFillBuffer (pattern of pattern_size bytes repeated over `bytes`). */
int    clrt_buf_fill(clrt_buf* buf, const void* h_pattern, size_t pattern_size, size_t bytes);
int    clrt_buf_upload(clrt_buf* buf, const void* h_src, size_t offset, size_t bytes);    /* enqueueWriteBuffer, blocking */
int    clrt_buf_download(const clrt_buf* buf, void* h_dst, size_t offset, size_t bytes);  /* CLBuffer::copyToHost, blocking */

/* Raw device-pointer copies: CLExecutionContext::enqueueWriteBuffer / enqueueReadBuffer
 * (R/src/OpenCLUtils/CLExecutionContext.cpp:386-470), blocking like the reference's. */
int    clrt_memcpy_htod(clrt_ctx* ctx, void* dptr_dst, const void* h_src, size_t bytes);
int    clrt_memcpy_dtoh(clrt_ctx* ctx, void* h_dst, const void* dptr_src, size_t bytes);

/* ---- building blocks ----------------------------------------------------------------------------------
 * BitonicSort::sort(cl_mem, size_t n)  R/Include/Algorithms/Sorting.h:54-85, R/src/CLRayTracerEngine/BitonicSort.cpp:118-189.
 *   Ascending sort of n (key,value) uint32 pairs by key, in place.  n is arbitrary here (the reference needs 2^k).
 *   mode CLRT_SORT_RADIX: onesweep LSD radix sort, STABLE (equal keys keep input order == the canonical
 *   tie order of SURVEY.md Q7 for every reference call site, whose values ascend with position).
 *   mode CLRT_SORT_BITONIC_COMPAT: reproduces the reference's unstable bitonic network permutation
 *   (n must be 2^k) for callers that need literal equality of tie order. */
#define CLRT_SORT_RADIX          0
#define CLRT_SORT_BITONIC_COMPAT 1
int clrt_sort_pairs_u32(clrt_ctx* ctx, void* dptr_pairs, size_t n, int mode);
/* PrefixSum::computePrefixSum(in,out,n)  R/Include/Algorithms/PrefixSum.h:48-83, R/src/CLRayTracerEngine/PrefixSum.cpp:94-133.
 *   INCLUSIVE scan of n uint32; in may equal out; any n. */
int clrt_scan_inclusive_u32(clrt_ctx* ctx, const void* dptr_in, void* dptr_out, size_t n);
/* The same scan for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that forbid
 *   cudaMalloc).  NO counterpart in the reference.  No context: `stream` is a
 *   cudaStream_t (NULL = the default stream) on the calling thread's CURRENT device; in / out are device pointers to n uint32,
 *   4-byte aligned, in may equal out.  The call makes no cudaMalloc / cudaFree / allocating library call: the tile descriptors
 *   (its only scratch memory) live in a static 128 KB device array inside the library, which limits n to
 *   CLRT_SCAN_NOALLOC_MAX_ELEMENTS — larger n fails with an error (use clrt_scan_inclusive_u32).  The first call on a device
 *   does the one-time set-up (make it outside stream capture).  Every call on a device shares that array: calls ordered on one
 *   stream or by events are fine, calls that OVERLAP on the GPU race and give undefined results.  Results are identical to
 *   clrt_scan_inclusive_u32. */
#define CLRT_SCAN_NOALLOC_MAX_ELEMENTS (1u << 28)
int clrt_scan_inclusive_u32_noalloc(void* stream, const void* dptr_in, void* dptr_out, size_t n);

/* ---- scene: Scene::loadToGPU (R/src/CLRayTracerEngine/Scene.cpp:451-456) ------------------------------
 * Uploads the packed v0.6 scene buffer (format: R/Include/CLData/SceneBufferParser.h) unchanged and
 * flattens it ON THE DEVICE into a 16-byte aligned clrt_triangle array used by every kernel. */
int      clrt_scene_upload(clrt_ctx* ctx, const void* h_scene, size_t bytes, clrt_scene** out);
int      clrt_scene_destroy(clrt_scene* scene);
uint64_t clrt_scene_triangle_count(const clrt_scene* scene);
void*    clrt_scene_dptr(const clrt_scene* scene);             /* Scene::getDeviceSceneData */
void*    clrt_scene_triangles_dptr(const clrt_scene* scene);   /* clrt_triangle[N]          */
int      clrt_scene_export_triangles(const clrt_scene* scene, void* h_out /* N*48 B */);

/* ---- acceleration structures: AccelerationStructureManager ------------------------------------------
 * BVHManager           R/Include/Algorithms/BVHManager.h:59-126,  R/src/CLRayTracerEngine/BVHManager.cpp
 * TwoLevelGridManager  R/Include/Algorithms/TwoLevelGridManager.h:61-178, R/src/CLRayTracerEngine/TwoLevelGridManager.cpp */
int clrt_bvh_create(clrt_ctx* ctx, const clrt_scene* scene, clrt_accel** out);                  /* ctor + initialize() */
int clrt_tlg_create(clrt_ctx* ctx, const clrt_scene* scene, float top_density, float leaf_density, clrt_accel** out);
/* KDTreeManager (ours: the reference's TODO 2.1, /root/reference/KNOWN_ISSUES_AND_TODO.txt:24-26 "Adding KD trees - Construction may be
 * based on method which is used to construct the BVH"): a third AccelerationStructureManager
 * (R/Include/Algorithms/AccelerationStructureManager.h:52-78) behind the same handle — clrt_accel_init_frame / _construct,
 * clrt_trace_primary / clrt_trace_rays, the frame entry points (three-call shadow pipeline).  Built like the LBVH (Morton codes of
 * the cells a triangle overlaps -> sort -> Karras radix tree over the occupied cells = a kd-tree of median planes); returns the true
 * closest hit, i.e. the BVH's contacts (openclraytracer_b200/csrc/kd.cu). */
int clrt_kdtree_create(clrt_ctx* ctx, const clrt_scene* scene, clrt_accel** out);
/* Inspection: out3 = {levels (2^levels cells per axis), (cell, triangle) pairs, leaves}; out_lo_cs6 = grid origin xyz, cell size xyz. */
int clrt_kd_info(clrt_accel* accel, uint32_t* out3, float* out_lo_cs6);
/* setTopLevelDensity / setLeafDensity (R/Include/Algorithms/TwoLevelGridManager.h:112-115): read at the next init_frame, like the
 * reference's calculateGridData (TwoLevelGridManager.cpp:584-600, called from initializeFrame). */
int clrt_tlg_set_densities(clrt_accel* accel, float top_density, float leaf_density);
/* getResolution() (TwoLevelGridManager.h:129, .cpp:566-578): top-level resolution for the current densities; out_xyz[3]. */
int clrt_tlg_resolution(clrt_accel* accel, uint32_t* out_xyz);
int clrt_accel_destroy(clrt_accel* accel);
int clrt_accel_init_frame(clrt_accel* accel);   /* initializeFrame()  BVHManager.cpp:122-150 / TwoLevelGridManager.cpp:160-201 */
int clrt_accel_construct(clrt_accel* accel);    /* construct()        BVHManager.cpp:156-228 / TwoLevelGridManager.cpp:207-469 */
/* generateContacts(Camera&)  BVHManager.cpp:234-281 -> kernel generateContacts (BVHKernels.cl:84-98).
 * Writes resX*resY clrt_contact to dptr_contacts (NULL: the accel's own primary-contacts buffer). */
int clrt_trace_primary(clrt_accel* accel, const clrt_camera* h_camera, void* dptr_contacts);
/* generateContacts(rays,contacts,rayCount)  BVHManager.cpp:292-329 -> kernel generateContacts2 (BVHKernels.cl:103-117). */
int clrt_trace_rays(clrt_accel* accel, const void* dptr_rays, void* dptr_contacts, uint32_t ray_count);
void* clrt_accel_primary_contacts_dptr(clrt_accel* accel);     /* getPrimaryContacts()->getCLMem() */
/* generateContacts(Camera&) for ONE image-space stripe (ours: the reference is single-device, CLExecutionContext.cpp:264-293):
 * tiles t = rank (mod nranks) of 32x8 pixels; contacts in stripe order (slot = local tile*256 + y*32 + x; slots outside the image
 * carry pixelIndex = UINT_MAX), clrt_frame_stripe_pixels() slots.  Feeds the generic-ray passes of a sharded frame. */
int clrt_trace_primary_stripe(clrt_accel* accel, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks, void* dptr_contacts);
/* Traversal policy.  FAST (default): near-child-first, t-pruned, persistent-thread kernel — same closest hit,
 * exact ties may pick another triangle.  REFERENCE_ORDER: the reference's A-before-B unpruned order
 * (R/Include/CLData/AccelerationStructs/BVH.h:344-408), bit-exact contacts; the parity vehicle.
 * On a Two-Level Grid: REFERENCE_ORDER (the grid's default) is the reference's walk literally (tlg_generate_contact,
 * R/Include/CLData/AccelerationStructs/TwoLevelGrid.h:557-678: entry t = min of the slab entries, the first leaf cell with any hit
 * ends the walk — its contact is not always the closest hit), bit-exact; FAST is the corrected two-level DDA that
 * returns the true closest hit, i.e. what the BVH returns (the reference's TODO, KNOWN_ISSUES_AND_TODO.txt:41-43). */
#define CLRT_TRACE_FAST            0
#define CLRT_TRACE_REFERENCE_ORDER 1
int clrt_accel_set_trace_mode(clrt_accel* accel, int mode);
/* Trace tree (ours; openclraytracer_b200/csrc/trace_tree.cu).  The reference-layout BVH is always built bit-exactly; when it
 * is a poor thing to traverse — equal-key chains (BVH.h:153-154,212-233) of 6+ leaves, or a scene box with more than 1/16 of an axis below zero so
 * that a Morton axis collapses (BVH.h:112-114 divides without subtracting min) — construct() also builds a plain LBVH over
 * the same triangles from a min-subtracted Morton code, and CLRT_TRACE_FAST walks that one.  Same closest hits; exact ties
 * go to the reference's winner (smallest position in the reference's sorted order).  policy: 0 auto (default), 1 always,
 * 2 never.  Takes effect at the next construct(). */
#define CLRT_TRACE_TREE_AUTO   0
#define CLRT_TRACE_TREE_ALWAYS 1
#define CLRT_TRACE_TREE_NEVER  2
int clrt_accel_set_trace_tree(clrt_accel* accel, int policy);
int clrt_accel_has_trace_tree(clrt_accel* accel, int* out);
/* Stage timings of the last construct() in ms (CUDA events on the context's stream). */
#define CLRT_STAGE_MORTON 0
#define CLRT_STAGE_SORT   1
#define CLRT_STAGE_EMIT   2
#define CLRT_STAGE_REFIT  3
#define CLRT_STAGE_TOTAL  4   /* everything construct() launched, the trace tree included */
#define CLRT_STAGE_TRACE_TREE 5 /* the separate traversal hierarchy (0 when none was built) */
/* Two-Level Grid stages (TwoLevelGridManager.cpp:207-469): count+scan, top pairs (scatter+sort+ranges), leaf cells,
 * leaf-pair count+scan, leaf pairs (exact scatter+sort+ranges). */
#define CLRT_STAGE_TLG_COUNT_TOP  8
#define CLRT_STAGE_TLG_TOP_PAIRS  9
#define CLRT_STAGE_TLG_LEAF_CELLS 10
#define CLRT_STAGE_TLG_LEAF_COUNT 11
#define CLRT_STAGE_TLG_LEAF_PAIRS 12
#define CLRT_STAGE_COUNT  16
int clrt_accel_stage_ms(clrt_accel* accel, int stage, float* out_ms);
int clrt_accel_enable_stage_timing(clrt_accel* accel, int on);
/* Traversal statistics mode (measurement only; the reference has none — SURVEY.md section 5): while on, the trace
 * kernels count rays traversed, inner-node visits and triangle tests.  set_stats zeroes the counters;
 * read_stats returns {rays, inner-node visits, triangle tests, leaf-box tests inside equal-key runs} accumulated since. */
/* Traversal stacks are bounded (the reference's is `stack[32]`, R/Include/CLData/AccelerationStructs/BVH.h:352, and overflows
 * silently).  Here a push that does not fit is counted on the device; this call waits for the context's stream and FAILS when a
 * launch since the last check dropped a subtree.  The blocking C++ / Python generateContacts call it after their wait. */
int clrt_accel_check(clrt_accel* accel);
int clrt_accel_set_stats(clrt_accel* accel, int on);
int clrt_accel_read_stats(clrt_accel* accel, uint64_t* h_out4);

/* Inspection of build products (the reference exposes the cl_mem objects; tests download them). */
int clrt_bvh_export_nodes(clrt_accel* accel, void* h_nodes /* (2N-1) * 48 B */);
int clrt_bvh_export_pairs(clrt_accel* accel, void* h_pairs /* N * 8 B, sorted (morton, leaf) */);
/* TLG products: GridData, TopLevelCell[cells], leafCellRanges[leafCells], sorted leaf pairs (real ones only). */
int clrt_tlg_sizes(clrt_accel* accel, uint32_t* top_cells, uint32_t* top_pairs, uint32_t* leaf_cells, uint32_t* leaf_pairs);
int clrt_tlg_export(clrt_accel* accel, clrt_grid_data* h_grid, void* h_top_cells, void* h_top_ranges,
                    void* h_top_pairs, void* h_leaf_ranges, void* h_leaf_pairs);

/* ---- ray generation / shading (host loop R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:298-316 ->
 *      cpuShadeBlinnPhong R/Include/CLData/Shading.h:74-166, generateRay R/Include/CLData/CLStructs.h:95-106) */
int clrt_generate_primary_rays(clrt_ctx* ctx, const clrt_camera* h_camera, void* dptr_rays /* resX*resY*48 B */);
/* One shadow ray per contact per light (scene lights, then the camera light).  Ray r = c*(L+1)+l goes from the
 * hit point (pushed 1e-3 along the geometric normal towards the light side) to light l; misses of the primary
 * pass produce a degenerate ray with idx = UINT_MAX that the trace kernels skip.  dptr_rays: n*(L+1)*48 B. */
int clrt_generate_shadow_rays(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera,
                              const void* dptr_contacts, uint32_t n, void* dptr_rays, uint32_t* out_rays_per_contact);
/* One cosine-weighted diffuse bounce ray per contact (incoherent secondary rays: BASELINE.json configs[3]; the reference's
 * TODO, KNOWN_ISSUES_AND_TODO.txt:27-33): ray i leaves the hit point of contact i (pushed 1e-4 * t along the normal turned
 * against the incoming ray) in a direction drawn from splitmix64(seed * 0x9E3779B97F4A7C15 + pixelIndex); misses give the
 * degenerate ray (idx = UINT_MAX, zero direction).  Trace them with clrt_trace_rays.  dptr_rays: n * 48 B. */
int clrt_generate_bounce_rays(clrt_ctx* ctx, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n, uint64_t seed,
                              void* dptr_rays);
/* Secondary rays of either kind from the contacts of ANY pass — the depth loop of the reference's TODO 2.2
 * (KNOWN_ISSUES_AND_TODO.txt: "generating rays from viewing rays-scene contacts, and execute ray tracing kernel in loop as many
 * times as specified by ray tracing depth"): rays_0 = contacts of clrt_trace_primary (dptr_in_rays == NULL, incoming ray from
 * h_camera), rays_k+1 = contacts of clrt_trace_rays(rays_k) with dptr_in_rays = rays_k (contact i belongs to ray i; h_camera may
 * be NULL; dptr_rays may alias dptr_in_rays).  The new ray keeps the pixel in idx; a miss or a dead incoming ray gives the
 * degenerate ray.  CLRT_RAYS_REFLECT: direction d - 2 (d.n) n about the normal turned against d (seed unused). */
#define CLRT_RAYS_DIFFUSE 0
#define CLRT_RAYS_REFLECT 1
int clrt_generate_secondary_rays(clrt_ctx* ctx, int kind, const clrt_camera* h_camera, const void* dptr_in_rays,
                                 const void* dptr_contacts, uint32_t n, uint64_t seed, void* dptr_rays);
/* Blinn-Phong of every contact -> RGBA8 (and/or float4) at the contact's slot.  dptr_shadow_contacts (may be
 * NULL) are the contacts of the shadow rays above: a light whose shadow ray hit before reaching it adds nothing. */
int clrt_shade(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera, const void* dptr_contacts,
               uint32_t n, const void* dptr_shadow_contacts, void* dptr_rgba8 /* may be NULL */,
               void* dptr_rgba_f32 /* may be NULL */);

/* Bounce frame (BASELINE.json configs[3]; ours — the reference has no secondary rays, KNOWN_ISSUES_AND_TODO.txt:27-33): Blinn-Phong of
 * contact i without shadows, each colour channel halved when the bounce ray of slot i (dptr_bounce_contacts[i], from
 * clrt_generate_bounce_rays + clrt_trace_rays) hit something: one-sample ambient occlusion -> RGBA8 at slot i. */
int clrt_shade_bounce(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n,
                      const void* dptr_bounce_contacts, void* dptr_rgba8);
/* One rank's frame of BASELINE.json configs[3] on a frame slot (ours; the reference has no secondary rays — KNOWN_ISSUES_AND_TODO.txt:27-30 —
 * and renders one frame at a time): clrt_trace_primary_stripe -> clrt_generate_bounce_rays -> clrt_trace_rays -> clrt_shade_bounce, enqueued
 * on `stream` (a cudaStream_t) with the slot's private work counter (slot 0 .. 15), so that frames can be in flight on several streams.
 * Buffers are the caller's, sized for the rank's stripe (clrt_frame_stripe_pixels): contacts and bounce contacts 32 B, rays 48 B, rgba8
 * 4 B per stripe slot; results identical to the four calls made one after the other.  BVH only. */
int clrt_render_bounce_stripe_on(clrt_accel* accel, void* stream, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                                 uint64_t seed, void* dptr_contacts, void* dptr_rays, void* dptr_bounce_contacts, void* dptr_rgba8);

/* Shadow pass + Blinn-Phong in ONE kernel over n contacts (BVH only): per contact and light an occlusion query along the
 * shadow ray clrt_generate_shadow_rays would build, then the shading of clrt_shade -> RGBA8.  Same pixels as the
 * three-call pipeline, without materialising rays and shadow contacts. */
int clrt_shadow_shade(clrt_accel* accel, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n, void* dptr_rgba8);

/* ---- frame: RTFrame() (OpenCLRayTracerDemo.cpp:271-326) as one call; the image-space shard entry -------
 * Renders the pixels of stripe `rank` of `nranks` (32x8-pixel tiles, tile t owned by rank t % nranks; inside the
 * GPU persistent CTAs pull tiles from an atomic counter) into dptr_rgba8 in STRIPE order (tile-major, 256 px per
 * tile).  shadows != 0 adds the shadow pass.  out_rays (host) receives the rays traced.  clrt_frame_untile
 * converts gathered stripes (rank-major) to raster order. */
int clrt_frame_stripe_pixels(uint32_t resX, uint32_t resY, uint32_t rank, uint32_t nranks, uint32_t* out_pixels, uint32_t* out_tiles);
int clrt_render_stripe(clrt_accel* accel, const clrt_camera* h_camera, int shadows, uint32_t rank, uint32_t nranks,
                       void* dptr_rgba8, uint64_t* out_rays);
/* The same on a caller-chosen cudaStream_t with frame slot `slot` (0..15): the slot's private work counter and scratch, so that
 * frames submitted on different streams overlap on the device — the tail of one frame's persistent kernels (a few long rays)
 * runs next to the head of the next frame's.  BVH with the fused shadow pass (or shadows == 0) only.  The caller orders reuse of
 * a slot and of dptr_rgba8 (stream order / events). */
int clrt_render_stripe_on(clrt_accel* accel, void* stream, int slot, const clrt_camera* h_camera, int shadows, uint32_t rank,
                          uint32_t nranks, void* dptr_rgba8, uint64_t* out_rays);
/* The frame entry points fuse the shadow pass with the shading on a BVH: on == 1 (default) two kernels — primary contacts, then
 * occlusion queries + Blinn-Phong over them; on == 2 ONE kernel for the whole stripe (primary ray, occlusion queries, shading; FAST
 * policy); on == 0 the three-call pipeline (clrt_generate_shadow_rays -> clrt_trace_rays -> clrt_shade).  Pixels are identical. */
int clrt_accel_set_fused_shadows(clrt_accel* accel, int on);
int clrt_frame_untile(clrt_ctx* ctx, const void* dptr_stripes, uint32_t resX, uint32_t resY, uint32_t nranks, void* dptr_raster_rgba8);
/* The same on a caller-chosen cudaStream_t (the frame pipeline runs gather + untile + read-back of frame k on a second
 * stream while frame k+1 is traced on the context's stream). */
int clrt_frame_untile_on(clrt_ctx* ctx, void* stream, const void* dptr_stripes, uint32_t resX, uint32_t resY, uint32_t nranks,
                         void* dptr_raster_rgba8);

/* ---- frame exchange over peer memory (one process per GPU of one box; new, the reference is single-device:
 * OpenCLRayTracerDemo.cpp:271-326).  Every process allocates its raster framebuffers and a block of flag words with
 * clrt_peer_alloc, exports them (64-byte cudaIpc handle, carried to the peers by any host channel) and opens its peers'.
 * A frame then needs no collective: each GPU stores the pixels of its own tiles into EVERY GPU's framebuffer over NVLink —
 * from the shading kernel (clrt_render_stripe_peers) or by one pass over its stripe buffer (clrt_frame_push) — raises a word on
 * every peer (clrt_peer_signal) and waits for the others' words (clrt_peer_wait).  All calls are asynchronous on `stream`. */
#define CLRT_PEER_HANDLE_BYTES 64
int clrt_peer_alloc(clrt_ctx* ctx, uint64_t bytes, void** out_dptr);            /* zeroed, exportable device memory */
int clrt_peer_free(clrt_ctx* ctx, void* dptr);
int clrt_peer_export(clrt_ctx* ctx, void* dptr, uint8_t* out_handle64);
int clrt_peer_open(clrt_ctx* ctx, const uint8_t* handle64, void** out_dptr);    /* another process's allocation */
int clrt_peer_close(clrt_ctx* ctx, void* dptr);
/* clrt_render_stripe_on with shadows, the pixels going to peer_rasters[0..nranks) (raster order, resX*resY RGBA8 each;
 * peer_rasters[rank] is this GPU's own) instead of a stripe buffer. */
int clrt_render_stripe_peers(clrt_accel* accel, void* stream, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                             void* const* peer_rasters, uint64_t* out_rays);
/* One call per frame and GPU: the whole protocol around clrt_render_stripe_peers.  peer_words[p] = 256 flag words of GPU p
 * (zeroed; ready at [slot*8 + writer], ack at [128 + slot*8 + writer]; 16 slots); seq = 1, 2, ... counts the frames of this slot.  Enqueued
 * on `stream`: wait until every GPU acked frame seq-1 of the slot (its framebuffer may be overwritten) -> trace + shade, pixels
 * to every framebuffer -> ready to all, wait for all -> h_rgba8 ? (copy this GPU's framebuffer to pinned host memory -> ack) : ack.
 * consumer_stream != NULL: everything from "wait for all" on is enqueued THERE (behind an event of this frame's "ready"), so the
 * trace stream of the slot never waits for the other GPUs' stripes nor for the read-back; the frame is complete on this GPU when
 * the consumer stream has passed.  Every GPU of the box must submit the same frames in the same order. */
int clrt_render_frame_peers(clrt_accel* accel, void* stream, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                            void* const* peer_rasters, void* const* peer_words, uint32_t seq, void* h_rgba8, uint64_t* out_rays,
                            void* consumer_stream);
/* `frames` (1..4) stripes (same rank / nranks, one camera each, one resolution) by ONE primary launch and ONE shadow + shade launch
 * into a local stripe-ordered buffer: frame f's stripe at dptr_rgba8 + f * (out_tiles of clrt_frame_stripe_pixels * 256) pixels; frame-slot semantics
 * of clrt_render_stripe_on.  BVH, FAST policy. */
int clrt_render_stripes_batch(clrt_accel* accel, void* stream, int slot, uint32_t frames, const clrt_camera* h_cameras, uint32_t rank,
                              uint32_t nranks, void* dptr_rgba8);
/* The same for `frames` (1..4) CONSECUTIVE frames — slots slot0 .. slot0+frames-1, cameras h_cameras[f] (one resolution), framebuffers
 * peer_rasters[f * 8 + gpu], sequence numbers seqs[f], read-back targets h_rgba8s[f] (array or entries may be NULL) — traced by ONE
 * primary launch and ONE shadow + shade launch: at 8 GPUs a stripe is an eighth of a frame and a launch of ~0.3 ms loses a sixth of
 * its time to its drain; four frames per launch make it a half-frame launch.  consumer_stream is required. */
int clrt_render_frames_peers(clrt_accel* accel, void* stream, int slot0, uint32_t frames, const clrt_camera* h_cameras, uint32_t rank,
                             uint32_t nranks, void* const* peer_rasters, void* const* peer_words, const uint32_t* seqs,
                             void* const* h_rgba8s, void* consumer_stream);
/* This GPU's stripe buffer (as written by clrt_render_stripe[_on]) -> raster position in every peer's framebuffer. */
int clrt_frame_push(clrt_ctx* ctx, void* stream, const void* dptr_stripe, uint32_t resX, uint32_t resY, uint32_t rank,
                    uint32_t nranks, void* const* peer_rasters);
/* peer_words[p][word] = value for p in 0..npeers, ordered behind everything enqueued on `stream` before (release, system scope). */
int clrt_peer_signal(clrt_ctx* ctx, void* stream, void* const* peer_words, uint32_t npeers, uint32_t word, uint32_t value);
/* Blocks `stream` (not the host) until words[first_word .. first_word+n) are all >= value (acquire, system scope).  After
 * timeout_ms (0: 30 s) the waiting kernel traps: the next synchronisation reports a CUDA error instead of hanging. */
int clrt_peer_wait(clrt_ctx* ctx, void* stream, const void* dptr_words, uint32_t first_word, uint32_t n, uint32_t value,
                   uint32_t timeout_ms);

/* Whole frame with HOST output: upload camera, trace, (shadow,) shade, download RGBA8 raster — the e2e path.
 * Camera.supersampilngFactor = s in 2..8 (R/Include/CLData/CLStructs.h:49; the reference sets 1 and never reads it — its TODO
 * 2.4) renders s*s rays per pixel: the frame of the camera (s*resX, s*resY, s*FOVDistance), box-filtered per channel with
 * rounding to resX x resY (alpha = coverage).  out_rays counts the sample rays.  0 and 1 mean one ray per pixel. */
int clrt_render_frame_host(clrt_accel* accel, const clrt_camera* h_camera, int shadows, void* h_rgba8, uint64_t* out_rays);

#ifdef __cplusplus
}
#endif
#endif
