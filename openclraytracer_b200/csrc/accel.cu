// accel.cu — the AccelerationStructureManager entry points of the C ABI
// (R/Include/Algorithms/AccelerationStructureManager.h:52-78; BVHManager.cpp / TwoLevelGridManager.cpp) and the
// per-frame flow of the Demo's RTFrame() (R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:271-326).
#include "common.cuh"
#include <vector>

int shade_launch_shadow_rays(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera& cam, const clrt_contact* contacts,
                             uint32_t n, clrt_ray* rays, unsigned long long* d_counter);
int shade_launch(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera& cam, const clrt_contact* contacts, uint32_t n,
                 const clrt_contact* shadow, uint32_t* rgba8, float4* rgba_f32, cudaStream_t stream);

int bvh_shadow_shade(clrt_accel* a, const clrt_camera& cam, const clrt_contact* contacts, uint32_t count, uint32_t* rgba8,
                     unsigned long long* d_ray_counter);
int bvh_frame_fused(clrt_accel* a, const clrt_camera& cam, uint32_t rank, uint32_t nranks, uint32_t* rgba8, unsigned long long* d_ray_counter);

static int accel_create(clrt_ctx* ctx, const clrt_scene* scene, AccelKind kind, clrt_accel** out) {
  if (!ctx || !scene || !out) CLRT_FAIL("accel create: NULL argument");
  if (scene->ctx != ctx) CLRT_FAIL("accel create: scene belongs to another context");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  clrt_accel* a = new clrt_accel();
  a->kind = kind; a->ctx = ctx; a->scene = scene;
  for (int i = 0; i <= CLRT_STAGE_COUNT; i++) CLRT_CUDA(cudaEventCreate(&a->ev[i]));
  CLRT_CUDA(cudaMalloc(&a->d_tile_counter, 64));
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 64, ctx->stream));
  CLRT_CUDA(cudaMalloc(&a->d_slot_counters, 64 * clrt_accel::FRAME_SLOTS));
  CLRT_CUDA(cudaMemsetAsync(a->d_slot_counters, 0, 64 * clrt_accel::FRAME_SLOTS, ctx->stream));
  *out = a;
  return 0;
}

extern "C" int clrt_bvh_create(clrt_ctx* ctx, const clrt_scene* scene, clrt_accel** out) {
  return accel_create(ctx, scene, ACCEL_BVH, out);
}
extern "C" int clrt_tlg_create(clrt_ctx* ctx, const clrt_scene* scene, float top_density, float leaf_density, clrt_accel** out) {
  if (!(top_density > 0.0f) || !(leaf_density > 0.0f)) CLRT_FAIL("clrt_tlg_create: densities must be positive");
  CLRT_TRY(accel_create(ctx, scene, ACCEL_TLG, out));
  (*out)->top_density = top_density; (*out)->leaf_density = leaf_density;
  (*out)->trace_mode = CLRT_TRACE_REFERENCE_ORDER;            // a grid answers like the reference's walk unless asked for the corrected one
  return 0;
}

// KDTreeManager (ours; the reference's TODO 2.1, KNOWN_ISSUES_AND_TODO.txt:24-26): a third AccelerationStructureManager, kd.cu
extern "C" int clrt_kdtree_create(clrt_ctx* ctx, const clrt_scene* scene, clrt_accel** out) {
  return accel_create(ctx, scene, ACCEL_KD, out);
}

extern "C" int clrt_accel_destroy(clrt_accel* a) {
  if (!a) return 0;
  cudaSetDevice(a->ctx->device);
  cudaStreamSynchronize(a->ctx->stream);
  if (a->kind == ACCEL_TLG) tlg_destroy(a);
  if (a->kind == ACCEL_KD) kd_destroy(a);
  if (a->d_nodes) cudaFree(a->d_nodes);
  if (a->d_pairs) cudaFree(a->d_pairs);
  if (a->d_pyr) cudaFree(a->d_pyr);
  if (a->d_tnodes) cudaFree(a->d_tnodes);
  if (a->d_refpos) cudaFree(a->d_refpos);
  trace_tree_destroy(a);
  if (a->d_tris_sorted) cudaFree(a->d_tris_sorted);
  if (a->d_contacts) cudaFree(a->d_contacts);
  if (a->d_frame_scratch) cudaFree(a->d_frame_scratch);
  if (a->d_fb) cudaFree(a->d_fb);
  if (a->d_tile_counter) cudaFree(a->d_tile_counter);
  if (a->d_slot_counters) cudaFree(a->d_slot_counters);
  for (int i = 0; i < clrt_accel::FRAME_SLOTS; i++) if (a->d_slot_scratch[i]) cudaFree(a->d_slot_scratch[i]);
  for (int i = 0; i <= CLRT_STAGE_COUNT; i++) if (a->ev[i]) cudaEventDestroy(a->ev[i]);
  for (int i = 0; i < clrt_accel::FRAME_SLOTS; i++) if (a->slot_event[i]) cudaEventDestroy(a->slot_event[i]);
  delete a;
  return 0;
}

extern "C" int clrt_accel_init_frame(clrt_accel* a) {
  if (!a) CLRT_FAIL("clrt_accel_init_frame: NULL accel");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  return a->kind == ACCEL_BVH ? bvh_init_frame(a) : a->kind == ACCEL_TLG ? tlg_init_frame(a) : kd_init_frame(a);
}
extern "C" int clrt_accel_construct(clrt_accel* a) {
  if (!a) CLRT_FAIL("clrt_accel_construct: NULL accel");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  return a->kind == ACCEL_BVH ? bvh_construct(a) : a->kind == ACCEL_TLG ? tlg_construct(a) : kd_construct(a);
}
extern "C" int clrt_accel_set_trace_mode(clrt_accel* a, int mode) {
  if (!a) CLRT_FAIL("clrt_accel_set_trace_mode: NULL accel");
  if (mode != CLRT_TRACE_FAST && mode != CLRT_TRACE_REFERENCE_ORDER) CLRT_FAIL("clrt_accel_set_trace_mode: unknown mode %d", mode);
  a->trace_mode = mode; return 0;
}
extern "C" int clrt_accel_set_fused_shadows(clrt_accel* a, int on) {
  if (!a) CLRT_FAIL("clrt_accel_set_fused_shadows: NULL accel");
  // 0: three kernels (primary, generic shadow rays, shade); 1: the default — two kernels (primary, shadow + shade); 2: one frame kernel
  a->unfused_shadows = on == 0; a->one_kernel_frame = on == 2; return 0;
}
extern "C" int clrt_accel_enable_stage_timing(clrt_accel* a, int on) {
  if (!a) CLRT_FAIL("clrt_accel_enable_stage_timing: NULL accel");
  a->timing = on != 0; return 0;
}
extern "C" int clrt_accel_set_stats(clrt_accel* a, int on) {
  if (!a) CLRT_FAIL("clrt_accel_set_stats: NULL accel");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter + 4, 0, 32, a->ctx->stream));
  a->stats = on != 0; return 0;
}
extern "C" int clrt_accel_read_stats(clrt_accel* a, uint64_t* out3) {
  if (!a || !out3) CLRT_FAIL("clrt_accel_read_stats: NULL argument");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  CLRT_CUDA(cudaMemcpyAsync(out3, a->d_tile_counter + 4, 32, cudaMemcpyDeviceToHost, a->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(a->ctx->stream));
  return 0;
}
extern "C" int clrt_accel_stage_ms(clrt_accel* a, int stage, float* out_ms) {
  if (!a || !out_ms) CLRT_FAIL("clrt_accel_stage_ms: NULL argument");
  if (stage < 0 || stage >= CLRT_STAGE_COUNT) CLRT_FAIL("clrt_accel_stage_ms: stage %d out of range", stage);
  *out_ms = a->stage_ms[stage]; return 0;
}

static int ensure_contacts(clrt_accel* a, size_t count) {
  const size_t bytes = count * sizeof(clrt_contact);
  if (bytes > a->contacts_cap) {
    CLRT_CUDA(cudaStreamSynchronize(a->ctx->stream));
    CLRT_TRY(clrt_grow((void**)&a->d_contacts, &a->contacts_cap, bytes));    // BVHManager.cpp:240-244 (resize)
  }
  return 0;
}

int accel_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks);
static int accel_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count) {
  return a->kind == ACCEL_BVH ? bvh_trace_rays(a, rays, out, count) : a->kind == ACCEL_TLG ? tlg_trace_rays(a, rays, out, count) : kd_trace_rays(a, rays, out, count);
}

extern "C" int clrt_trace_primary(clrt_accel* a, const clrt_camera* h_camera, void* dptr_contacts) {
  if (!a || !h_camera) CLRT_FAIL("clrt_trace_primary: NULL argument");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  const uint64_t pixels = (uint64_t)h_camera->resX * h_camera->resY;
  if (pixels > 0x7FFFFFFFull) CLRT_FAIL("clrt_trace_primary: resolution too large");
  clrt_contact* out = (clrt_contact*)dptr_contacts;
  if (!out) { CLRT_TRY(ensure_contacts(a, pixels)); out = a->d_contacts; }
  return accel_trace_primary(a, *h_camera, out, 0, 0);
}

// Stripe `rank` of `nranks` of the camera's primary rays (image-space sharding): contacts in STRIPE order — slot =
// local tile * 256 + y * 32 + x, slots outside the image get pixelIndex = UINT_MAX — for clrt_frame_stripe_pixels() slots.
extern "C" int clrt_trace_primary_stripe(clrt_accel* a, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks, void* dptr_contacts) {
  if (!a || !h_camera || !dptr_contacts) CLRT_FAIL("clrt_trace_primary_stripe: NULL argument");
  if (!nranks || rank >= nranks) CLRT_FAIL("clrt_trace_primary_stripe: rank %u of %u", rank, nranks);
  if ((uintptr_t)dptr_contacts & 15) CLRT_FAIL("clrt_trace_primary_stripe: the contact buffer must be 16-byte aligned");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  if ((uint64_t)h_camera->resX * h_camera->resY > 0x7FFFFFFFull) CLRT_FAIL("clrt_trace_primary_stripe: resolution too large");
  return accel_trace_primary(a, *h_camera, (clrt_contact*)dptr_contacts, rank, nranks);
}

extern "C" void* clrt_accel_primary_contacts_dptr(clrt_accel* a) { return a ? a->d_contacts : nullptr; }

extern "C" int clrt_trace_rays(clrt_accel* a, const void* dptr_rays, void* dptr_contacts, uint32_t ray_count) {
  if (!a || (ray_count && (!dptr_rays || !dptr_contacts))) CLRT_FAIL("clrt_trace_rays: NULL argument");
  if (((uintptr_t)dptr_rays & 15) || ((uintptr_t)dptr_contacts & 15)) CLRT_FAIL("clrt_trace_rays: buffers must be 16-byte aligned");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  return accel_trace_rays(a, (const clrt_ray*)dptr_rays, (clrt_contact*)dptr_contacts, ray_count);
}

// configs[3] frame of one rank on a frame slot: primary contacts of the rank's stripe -> one diffuse bounce ray per hit -> the generic-ray
// walk -> shading, all enqueued on `stream` with the slot's private work counter, so that several frames can be in flight on several
// streams (frame.BounceRenderer).  The four steps are the public calls clrt_trace_primary_stripe, clrt_generate_bounce_rays,
// clrt_trace_rays, clrt_shade_bounce, unchanged; only where they are enqueued differs.  Ours (the reference renders one frame at a time).
extern "C" int clrt_render_bounce_stripe_on(clrt_accel* a, void* stream_, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                                            uint64_t seed, void* dptr_contacts, void* dptr_rays, void* dptr_bounce_contacts, void* dptr_rgba8) {
  if (!a || !h_camera || !dptr_contacts || !dptr_rays || !dptr_bounce_contacts || !dptr_rgba8) CLRT_FAIL("clrt_render_bounce_stripe_on: NULL argument");
  if (a->kind != ACCEL_BVH) CLRT_FAIL("clrt_render_bounce_stripe_on: frame slots exist for the BVH only");
  if (slot < 0 || slot >= clrt_accel::FRAME_SLOTS) CLRT_FAIL("clrt_render_bounce_stripe_on: slot %d of %d", slot, clrt_accel::FRAME_SLOTS);
  if (!nranks || rank >= nranks) CLRT_FAIL("clrt_render_bounce_stripe_on: rank %u of %u", rank, nranks);
  clrt_ctx* ctx = a->ctx;
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (slot + 1 > a->slots_seen) a->slots_seen = slot + 1;
  struct Env {                                    // lends the stream to everything the four calls enqueue, and the slot's counter to the trace launches
    clrt_accel* a; cudaStream_t saved;
    Env(clrt_accel* a_, cudaStream_t s, uint32_t* c) : a(a_), saved(a_->ctx->stream) { a->ctx->stream = s; a->launch_stream = s; a->launch_counter = c; }
    ~Env() { a->ctx->stream = saved; a->launch_stream = nullptr; a->launch_counter = nullptr; }
  } env(a, (cudaStream_t)stream_, a->d_slot_counters + 16 * slot);
  const TileGrid g = make_tile_grid(h_camera->resX, h_camera->resY);
  const uint32_t units = stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  if (!units) return 0;
  CLRT_TRY(clrt_trace_primary_stripe(a, h_camera, rank, nranks, dptr_contacts));
  CLRT_TRY(clrt_generate_bounce_rays(ctx, h_camera, dptr_contacts, units, seed, dptr_rays));
  CLRT_TRY(clrt_trace_rays(a, dptr_rays, dptr_bounce_contacts, units));
  CLRT_TRY(clrt_shade_bounce(ctx, a->scene, h_camera, dptr_contacts, units, dptr_bounce_contacts, dptr_rgba8));
  return 0;
}

// Shadow pass fused with the shading over an arbitrary contact buffer (BVH only).
extern "C" int clrt_shadow_shade(clrt_accel* a, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n, void* dptr_rgba8) {
  if (!a || !h_camera || (n && (!dptr_contacts || !dptr_rgba8))) CLRT_FAIL("clrt_shadow_shade: NULL argument");
  if (a->kind != ACCEL_BVH) CLRT_FAIL("clrt_shadow_shade: the fused shadow pass exists for the BVH only");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  return bvh_shadow_shade(a, *h_camera, (const clrt_contact*)dptr_contacts, n, (uint32_t*)dptr_rgba8, nullptr);
}

// ---- frame ---------------------------------------------------------------------------------------------
extern "C" int clrt_frame_stripe_pixels(uint32_t resX, uint32_t resY, uint32_t rank, uint32_t nranks, uint32_t* out_pixels, uint32_t* out_tiles) {
  if (!nranks || rank >= nranks) CLRT_FAIL("clrt_frame_stripe_pixels: rank %u of %u", rank, nranks);
  const TileGrid g = make_tile_grid(resX, resY);
  if (out_pixels) *out_pixels = ((g.tiles + nranks - 1) / nranks) * CLRT_TILE_PIXELS;     // padded stripe capacity
  if (out_tiles) *out_tiles = stripe_tile_count(g.tiles, rank, nranks);
  return 0;
}

// scratch layout of one stripe: contacts[slots] | shadow rays[slots*(L+1)] | shadow contacts[slots*(L+1)] | ray counter
// slot < 0: the context's stream, the accel's own scratch and counter.  slot >= 0: the caller's stream with the slot's private
// scratch and counter, so that frames submitted on different streams overlap on the device (the tail of one frame's persistent
// kernels — a few long rays — runs next to the head of the next frame's).
static int render_stripe_impl(clrt_accel* a, const clrt_camera* h_camera, int shadows, uint32_t rank, uint32_t nranks,
                              void* dptr_rgba8, uint64_t* out_rays, cudaStream_t stream, int slot) {
  if (!a || !h_camera || !dptr_rgba8) CLRT_FAIL("clrt_render_stripe: NULL argument");
  if (!nranks || rank >= nranks) CLRT_FAIL("clrt_render_stripe: rank %u of %u", rank, nranks);
  clrt_ctx* ctx = a->ctx;
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (slot >= clrt_accel::FRAME_SLOTS) CLRT_FAIL("clrt_render_stripe_on: slot %d of %d", slot, clrt_accel::FRAME_SLOTS);
  if (slot >= 0 && a->kind != ACCEL_BVH) CLRT_FAIL("clrt_render_stripe_on: frame slots exist for the BVH only");
  if (slot < 0) stream = ctx->stream;
  if (slot + 1 > a->slots_seen) a->slots_seen = slot + 1;
  void** scratch_p = slot < 0 ? &a->d_frame_scratch : &a->d_slot_scratch[slot];
  size_t* scratch_cap = slot < 0 ? &a->frame_scratch_cap : &a->slot_scratch_cap[slot];
  struct LaunchEnv {                              // lends the slot's stream + counter to the trace launches of this call
    clrt_accel* a;
    LaunchEnv(clrt_accel* a_, cudaStream_t s, uint32_t* c) : a(a_) { a->launch_stream = s; a->launch_counter = c; }
    ~LaunchEnv() { a->launch_stream = nullptr; a->launch_counter = nullptr; }
  } env(a, slot < 0 ? nullptr : stream, slot < 0 ? nullptr : a->d_slot_counters + 16 * slot);
  const clrt_camera cam = *h_camera;
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const uint32_t tiles = stripe_tile_count(g.tiles, rank, nranks);
  const uint32_t slots = tiles * CLRT_TILE_PIXELS;
  const uint32_t L1 = a->scene->n_lights + 1;
  // BVH: the shadow pass is fused with the shading (occlusion queries, no ray / shadow-contact buffers); the grid keeps
  // the three-kernel pipeline through the generic-ray entry point.
  const bool fused = shadows && a->kind == ACCEL_BVH && L1 <= 32 && !a->unfused_shadows;
  const size_t contacts_bytes = (size_t)slots * sizeof(clrt_contact);
  const size_t srays_bytes = (shadows && !fused) ? (size_t)slots * L1 * sizeof(clrt_ray) : 0;
  const size_t scont_bytes = (shadows && !fused) ? (size_t)slots * L1 * sizeof(clrt_contact) : 0;
  const size_t total = contacts_bytes + srays_bytes + scont_bytes + 64;
  if (total > *scratch_cap) {
    CLRT_CUDA(cudaStreamSynchronize(stream));
    CLRT_TRY(clrt_grow(scratch_p, scratch_cap, total));
  }
  uint8_t* base = (uint8_t*)*scratch_p;
  clrt_contact* contacts = (clrt_contact*)base;
  clrt_ray* srays = (clrt_ray*)(base + contacts_bytes);
  clrt_contact* scont = (clrt_contact*)(base + contacts_bytes + srays_bytes);
  unsigned long long* counter = (unsigned long long*)(base + contacts_bytes + srays_bytes + scont_bytes);
  // Two launches by default: primary contacts, then shadow + shade over them.  clrt_accel_set_fused_shadows(2) runs the whole stripe
  // as ONE launch (FAST policy only): one drain instead of two, no contacts through memory — but warps that mix closest-hit and
  // occlusion rays lose coherence: measured on Soup-1M at 1080p, one GPU, 2.45 ms per frame against 2.31 (profiles/r02_experiments.md).
  const bool one_kernel = fused && a->trace_mode == CLRT_TRACE_FAST && a->one_kernel_frame;
  if (slots && one_kernel) {
    if (out_rays) CLRT_CUDA(cudaMemsetAsync(counter, 0, 8, stream));
    CLRT_TRY(bvh_frame_fused(a, cam, rank, nranks, (uint32_t*)dptr_rgba8, out_rays ? counter : nullptr));
  } else if (slots) {
    CLRT_TRY(accel_trace_primary(a, cam, contacts, rank, nranks));
    if (fused) {
      CLRT_CUDA(cudaMemsetAsync(counter, 0, 8, stream));
      CLRT_TRY(bvh_shadow_shade(a, cam, contacts, slots, (uint32_t*)dptr_rgba8, out_rays ? counter : nullptr));
    } else if (shadows) {
      CLRT_CUDA(cudaMemsetAsync(counter, 0, 8, ctx->stream));
      CLRT_TRY(shade_launch_shadow_rays(ctx, a->scene, cam, contacts, slots, srays, out_rays ? counter : nullptr));
      CLRT_TRY(accel_trace_rays(a, srays, scont, slots * L1));
    }
    if (!fused && shadows && slot >= 0) CLRT_FAIL("clrt_render_stripe_on: frame slots need the fused shadow pass");
    if (!fused) CLRT_TRY(shade_launch(ctx, a->scene, cam, contacts, slots, shadows ? scont : nullptr, (uint32_t*)dptr_rgba8, nullptr, stream));
  }
  if (out_rays) {
    // primary rays = pixels of this stripe that lie inside the image; shadow rays = live rays counted on the device
    uint64_t primary = 0;
    for (uint32_t lt = 0; lt < tiles; lt++) {
      const uint32_t gt = rank + lt * nranks;
      const uint32_t x0 = (gt % g.tilesX) * CLRT_TILE_W, y0 = (gt / g.tilesX) * CLRT_TILE_H;
      const uint32_t w = x0 + CLRT_TILE_W <= cam.resX ? CLRT_TILE_W : cam.resX - x0;
      const uint32_t h = y0 + CLRT_TILE_H <= cam.resY ? CLRT_TILE_H : cam.resY - y0;
      primary += (uint64_t)w * h;
    }
    unsigned long long srays_count = 0;
    if (shadows && slots) CLRT_CUDA(cudaMemcpyAsync(&srays_count, counter, 8, cudaMemcpyDeviceToHost, stream));
    CLRT_CUDA(cudaStreamSynchronize(stream));
    *out_rays = primary + srays_count;
  }
  return 0;
}

extern "C" int clrt_render_stripe(clrt_accel* a, const clrt_camera* h_camera, int shadows, uint32_t rank, uint32_t nranks,
                                  void* dptr_rgba8, uint64_t* out_rays) {
  return render_stripe_impl(a, h_camera, shadows, rank, nranks, dptr_rgba8, out_rays, nullptr, -1);
}
extern "C" int clrt_render_stripe_on(clrt_accel* a, void* stream, int slot, const clrt_camera* h_camera, int shadows, uint32_t rank,
                                     uint32_t nranks, void* dptr_rgba8, uint64_t* out_rays) {
  if (slot < 0) CLRT_FAIL("clrt_render_stripe_on: slot %d", slot);
  return render_stripe_impl(a, h_camera, shadows, rank, nranks, dptr_rgba8, out_rays, (cudaStream_t)stream, slot);
}

// The fused exchange: the shading kernel stores each pixel of this GPU's tiles into the raster-ordered framebuffer of every GPU of
// the box (peer.cu); no stripe buffer, no gather, no untile.
extern "C" int clrt_render_stripe_peers(clrt_accel* a, void* stream, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                                        void* const* peer_rasters, uint64_t* out_rays);
extern "C" int clrt_render_stripe_peers(clrt_accel* a, void* stream, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                                        void* const* peer_rasters, uint64_t* out_rays) {
  if (!a || !peer_rasters || !nranks || nranks > 8 || rank >= nranks) CLRT_FAIL("clrt_render_stripe_peers: bad argument");
  if (slot < 0) CLRT_FAIL("clrt_render_stripe_peers: slot %d", slot);
  if (a->kind != ACCEL_BVH || a->scene->n_lights + 1 > 32 || a->unfused_shadows)
    CLRT_FAIL("clrt_render_stripe_peers: needs the BVH with the fused shadow pass");
  struct Targets {
    clrt_accel* a;
    Targets(clrt_accel* a_, void* const* p, uint32_t n) : a(a_) { a->launch_targets.n = n; for (uint32_t i = 0; i < 8; i++) a->launch_targets.p[i] = i < n ? (uint32_t*)p[i] : nullptr; }
    ~Targets() { a->launch_targets.n = 0; }
  } targets(a, peer_rasters, nranks);
  for (uint32_t i = 0; i < nranks; i++) if (!peer_rasters[i]) CLRT_FAIL("clrt_render_stripe_peers: peer %u is NULL", i);
  return render_stripe_impl(a, h_camera, 1, rank, nranks, peer_rasters[rank], out_rays, (cudaStream_t)stream, slot);
}

int peer_launch_wait(cudaStream_t stream, const uint32_t* words, uint32_t n, uint32_t value);
int peer_launch_exchange(cudaStream_t stream, const PeerTargets& flags, const uint32_t* own, uint32_t ready_base, uint32_t ack_base,
                         uint32_t rank, uint32_t seq);
int peer_launch_signal(cudaStream_t stream, const PeerTargets& flags, uint32_t word, uint32_t value);

// One call per frame and GPU.  Flag words of a GPU (peer_words[p], 256 words): ready at [slot*8 + writer], ack at [128 + slot*8 +
// writer]; seq = 1, 2, ... per slot.  Enqueues on `stream`: wait until every GPU acked seq-1 of this slot (its framebuffer is
// free) -> trace + shade with the pixels going to every framebuffer -> ready to all, wait for all -> (copy to host ->) ack to all.
extern "C" int clrt_render_frame_peers(clrt_accel* a, void* stream_, int slot, const clrt_camera* h_camera, uint32_t rank, uint32_t nranks,
                                       void* const* peer_rasters, void* const* peer_words, uint32_t seq, void* h_rgba8, uint64_t* out_rays,
                                       void* consumer_stream_) {
  if (!a || !h_camera || !peer_rasters || !peer_words || !nranks || nranks > 8 || rank >= nranks || !seq)
    CLRT_FAIL("clrt_render_frame_peers: bad argument");
  if (slot < 0 || slot >= clrt_accel::FRAME_SLOTS) CLRT_FAIL("clrt_render_frame_peers: slot %d of %d", slot, clrt_accel::FRAME_SLOTS);
  cudaStream_t stream = (cudaStream_t)stream_, consumer = (cudaStream_t)consumer_stream_;
  PeerTargets flags; flags.n = nranks;
  for (uint32_t i = 0; i < 8; i++) flags.p[i] = i < nranks ? (uint32_t*)peer_words[i] : nullptr;
  for (uint32_t i = 0; i < nranks; i++) if (!flags.p[i]) CLRT_FAIL("clrt_render_frame_peers: peer %u is NULL", i);
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  const uint32_t* own = flags.p[rank];
  const uint32_t ready = (uint32_t)slot * 8u, ack = 128u + (uint32_t)slot * 8u;
  if (seq > 1) CLRT_TRY(peer_launch_wait(stream, own + ack, nranks, seq - 1));
  CLRT_TRY(clrt_render_stripe_peers(a, stream_, slot, h_camera, rank, nranks, peer_rasters, out_rays));
  const size_t bytes = (size_t)h_camera->resX * h_camera->resY * 4;
  if (!consumer) {
    CLRT_TRY(peer_launch_exchange(stream, flags, own, ready, h_rgba8 ? ~0u : ack, rank, seq));
    if (h_rgba8) {
      CLRT_CUDA(cudaMemcpyAsync(h_rgba8, peer_rasters[rank], bytes, cudaMemcpyDeviceToHost, stream));
      CLRT_TRY(peer_launch_signal(stream, flags, ack + rank, seq));
    }
    return 0;
  }
  // Trace stream: ready to all, nothing else — it goes on with the slot's next frame as soon as the acks of this one arrive.
  // Consumer stream: behind OUR ready (an event, so that its spinning wait can never sit in front of the kernels it waits for when
  // streams share a hardware queue), wait for everybody's ready -> read-back -> ack to all.
  CLRT_TRY(peer_launch_signal(stream, flags, ready + rank, seq));
  if (!a->slot_event[slot]) CLRT_CUDA(cudaEventCreateWithFlags(&a->slot_event[slot], cudaEventDisableTiming));
  CLRT_CUDA(cudaEventRecord(a->slot_event[slot], stream));
  CLRT_CUDA(cudaStreamWaitEvent(consumer, a->slot_event[slot], 0));
  CLRT_TRY(peer_launch_wait(consumer, own + ready, nranks, seq));
  if (h_rgba8) CLRT_CUDA(cudaMemcpyAsync(h_rgba8, peer_rasters[rank], bytes, cudaMemcpyDeviceToHost, consumer));
  CLRT_TRY(peer_launch_signal(consumer, flags, ack + rank, seq));
  return 0;
}

int bvh_frames_batch(clrt_accel* a, const clrt_camera* cams, uint32_t frames, uint32_t rank, uint32_t nranks, clrt_contact* contacts,
                     uint32_t* rgba8, uint32_t* const* targets, uint32_t n_targets, unsigned long long* d_ray_counter);

// `frames` (1..4) stripes (same rank / nranks, one camera each) by ONE primary launch and ONE shadow + shade launch into a local
// stripe-ordered buffer: frame f's stripe at dptr_rgba8 + f * (this rank's tiles * 256) pixels.  Frame-slot semantics of
// clrt_render_stripe_on (the slot's counter and scratch, the caller's stream).
extern "C" int clrt_render_stripes_batch(clrt_accel* a, void* stream_, int slot, uint32_t frames, const clrt_camera* h_cameras, uint32_t rank,
                                         uint32_t nranks, void* dptr_rgba8) {
  if (!a || !h_cameras || !dptr_rgba8 || !nranks || rank >= nranks) CLRT_FAIL("clrt_render_stripes_batch: bad argument");
  if (!frames || frames > 4 || slot < 0 || slot >= clrt_accel::FRAME_SLOTS) CLRT_FAIL("clrt_render_stripes_batch: frames %u, slot %d", frames, slot);
  if (a->kind != ACCEL_BVH || a->scene->n_lights + 1 > 32 || a->unfused_shadows || a->trace_mode != CLRT_TRACE_FAST)
    CLRT_FAIL("clrt_render_stripes_batch: needs the BVH with the fused shadow pass and the FAST policy");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  cudaStream_t stream = (cudaStream_t)stream_;
  if (slot + 1 > a->slots_seen) a->slots_seen = slot + 1;
  const TileGrid g = make_tile_grid(h_cameras[0].resX, h_cameras[0].resY);
  const size_t units = (size_t)stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  const size_t need = units * frames * sizeof(clrt_contact) + 64;
  if (need > a->slot_scratch_cap[slot]) {
    CLRT_CUDA(cudaStreamSynchronize(stream));
    CLRT_TRY(clrt_grow(&a->d_slot_scratch[slot], &a->slot_scratch_cap[slot], need));
  }
  struct LaunchEnv {
    clrt_accel* a;
    LaunchEnv(clrt_accel* a_, cudaStream_t s, uint32_t* c, int batch) : a(a_) { a->launch_stream = s; a->launch_counter = c; a->launch_batch = batch; }
    ~LaunchEnv() { a->launch_stream = nullptr; a->launch_counter = nullptr; a->launch_batch = 1; }
  } env(a, stream, a->d_slot_counters + 16 * slot, 1);
  return bvh_frames_batch(a, h_cameras, frames, rank, nranks, (clrt_contact*)a->d_slot_scratch[slot], (uint32_t*)dptr_rgba8, nullptr, 0, nullptr);
}

// `frames` (1..4) consecutive frames of the fused exchange as ONE pair of launches: slots slot0 .. slot0+frames-1, their cameras
// h_cameras[f], their framebuffers peer_rasters[f * 8 + gpu] and sequence numbers seqs[f].  Protocol per frame as in
// clrt_render_frame_peers with a consumer stream (required); what changes is that the stripes of all the frames are traced by one
// primary launch and one shadow + shade launch, so that at 8 GPUs a launch carries half a frame's work instead of an eighth and
// loses a quarter as much to its drain.  h_rgba8s (may be NULL) = frames host buffers for the read-back.
extern "C" int clrt_render_frames_peers(clrt_accel* a, void* stream_, int slot0, uint32_t frames, const clrt_camera* h_cameras, uint32_t rank,
                                        uint32_t nranks, void* const* peer_rasters, void* const* peer_words, const uint32_t* seqs,
                                        void* const* h_rgba8s, void* consumer_stream_) {
  if (!a || !h_cameras || !peer_rasters || !peer_words || !seqs || !consumer_stream_ || !nranks || nranks > 8 || rank >= nranks)
    CLRT_FAIL("clrt_render_frames_peers: bad argument");
  if (!frames || frames > 4 || slot0 < 0 || slot0 + (int)frames > clrt_accel::FRAME_SLOTS)
    CLRT_FAIL("clrt_render_frames_peers: frames %u from slot %d (of %d)", frames, slot0, clrt_accel::FRAME_SLOTS);
  if (a->kind != ACCEL_BVH || a->scene->n_lights + 1 > 32 || a->unfused_shadows || a->trace_mode != CLRT_TRACE_FAST)
    CLRT_FAIL("clrt_render_frames_peers: needs the BVH with the fused shadow pass and the FAST policy");
  cudaStream_t stream = (cudaStream_t)stream_, consumer = (cudaStream_t)consumer_stream_;
  PeerTargets flags; flags.n = nranks;
  for (uint32_t i = 0; i < 8; i++) flags.p[i] = i < nranks ? (uint32_t*)peer_words[i] : nullptr;
  for (uint32_t i = 0; i < nranks; i++) if (!flags.p[i]) CLRT_FAIL("clrt_render_frames_peers: peer %u is NULL", i);
  uint32_t* targets[4 * 8] = {nullptr};
  for (uint32_t f = 0; f < frames; f++) {
    if (!seqs[f]) CLRT_FAIL("clrt_render_frames_peers: sequence numbers start at 1");
    for (uint32_t k = 0; k < nranks; k++) {
      targets[f * 8 + k] = (uint32_t*)peer_rasters[f * 8 + k];
      if (!targets[f * 8 + k]) CLRT_FAIL("clrt_render_frames_peers: framebuffer %u of frame %u is NULL", k, f);
    }
  }
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  if (slot0 + (int)frames > a->slots_seen) a->slots_seen = slot0 + (int)frames;
  const uint32_t* own = flags.p[rank];
  for (uint32_t f = 0; f < frames; f++)
    if (seqs[f] > 1) CLRT_TRY(peer_launch_wait(stream, own + 128u + (uint32_t)(slot0 + f) * 8u, nranks, seqs[f] - 1));
  // contacts of all the frames in slot0's scratch
  const TileGrid g = make_tile_grid(h_cameras[0].resX, h_cameras[0].resY);
  const size_t units = (size_t)stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  const size_t need = units * frames * sizeof(clrt_contact) + 64;
  if (need > a->slot_scratch_cap[slot0]) {
    CLRT_CUDA(cudaStreamSynchronize(stream));
    CLRT_TRY(clrt_grow(&a->d_slot_scratch[slot0], &a->slot_scratch_cap[slot0], need));
  }
  {
    struct LaunchEnv {
      clrt_accel* a;
      LaunchEnv(clrt_accel* a_, cudaStream_t s, uint32_t* c, int batch) : a(a_) { a->launch_stream = s; a->launch_counter = c; a->launch_batch = batch; }
      ~LaunchEnv() { a->launch_stream = nullptr; a->launch_counter = nullptr; a->launch_batch = 1; }
    } env(a, stream, a->d_slot_counters + 16 * slot0, (int)frames);
    CLRT_TRY(bvh_frames_batch(a, h_cameras, frames, rank, nranks, (clrt_contact*)a->d_slot_scratch[slot0], nullptr, targets, nranks, nullptr));
  }
  const size_t bytes = (size_t)h_cameras[0].resX * h_cameras[0].resY * 4;
  for (uint32_t f = 0; f < frames; f++) CLRT_TRY(peer_launch_signal(stream, flags, (uint32_t)(slot0 + f) * 8u + rank, seqs[f]));
  if (!a->slot_event[slot0]) CLRT_CUDA(cudaEventCreateWithFlags(&a->slot_event[slot0], cudaEventDisableTiming));
  CLRT_CUDA(cudaEventRecord(a->slot_event[slot0], stream));
  CLRT_CUDA(cudaStreamWaitEvent(consumer, a->slot_event[slot0], 0));
  for (uint32_t f = 0; f < frames; f++) {
    CLRT_TRY(peer_launch_wait(consumer, own + (uint32_t)(slot0 + f) * 8u, nranks, seqs[f]));
    if (h_rgba8s && h_rgba8s[f]) CLRT_CUDA(cudaMemcpyAsync(h_rgba8s[f], targets[f * 8 + rank], bytes, cudaMemcpyDeviceToHost, consumer));
    CLRT_TRY(peer_launch_signal(consumer, flags, 128u + (uint32_t)(slot0 + f) * 8u + rank, seqs[f]));
  }
  return 0;
}

int shade_launch_downsample(clrt_ctx* ctx, const uint32_t* hi, uint32_t resX, uint32_t resY, uint32_t s, uint32_t* lo);

// Camera.supersampilngFactor (CLStructs.h:49; set to 1 and never read by the reference — its TODO 2.4) = s: the frame of the camera
// (s*resX, s*resY, s*FOVDistance) — s*s rays per pixel on a regular grid, same field of view — box-filtered to resX x resY.
extern "C" int clrt_render_frame_host(clrt_accel* a, const clrt_camera* h_camera, int shadows, void* h_rgba8, uint64_t* out_rays) {
  if (!a || !h_camera || !h_rgba8) CLRT_FAIL("clrt_render_frame_host: NULL argument");
  clrt_ctx* ctx = a->ctx;
  CLRT_CUDA(cudaSetDevice(ctx->device));
  const uint32_t ss = h_camera->supersampilngFactor > 1 ? h_camera->supersampilngFactor : 1;
  if (ss > 8) CLRT_FAIL("clrt_render_frame_host: supersampilngFactor %u (1..8)", ss);
  clrt_camera cam = *h_camera;
  cam.resX *= ss; cam.resY *= ss; cam.FOVDistance *= (float)ss; cam.supersampilngFactor = 1;
  if ((uint64_t)cam.resX * cam.resY > 0x7FFFFFFFull) CLRT_FAIL("clrt_render_frame_host: %u x %u samples", cam.resX, cam.resY);
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const size_t stripe_bytes = (size_t)g.tiles * CLRT_TILE_PIXELS * 4, raster_bytes = (size_t)cam.resX * cam.resY * 4;
  const size_t out_bytes = (size_t)h_camera->resX * h_camera->resY * 4;
  // framebuffers owned by the accel: [stripe | raster]; the filtered frame reuses the stripe area (free once untiled)
  if (stripe_bytes + raster_bytes + 256 > a->fb_cap) {
    CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
    CLRT_TRY(clrt_grow(&a->d_fb, &a->fb_cap, stripe_bytes + raster_bytes + 256));
  }
  uint8_t* stripe = (uint8_t*)a->d_fb; uint8_t* raster = stripe + ((stripe_bytes + 255) & ~(size_t)255);
  CLRT_TRY(clrt_render_stripe(a, &cam, shadows, 0, 1, stripe, nullptr));
  CLRT_TRY(clrt_frame_untile(ctx, stripe, cam.resX, cam.resY, 1, raster));
  const uint8_t* result = raster;
  if (ss > 1) {
    CLRT_TRY(shade_launch_downsample(ctx, (const uint32_t*)raster, h_camera->resX, h_camera->resY, ss, (uint32_t*)stripe));
    result = stripe;
  }
  CLRT_CUDA(cudaMemcpyAsync(h_rgba8, result, out_bytes, cudaMemcpyDeviceToHost, ctx->stream));
  CLRT_TRY(clrt_accel_check(a));                               // waits for the stream; a traversal-stack overflow fails the frame
  if (out_rays) {
    // counted without re-rendering: primary = all samples; shadow = (L+1) per hit sample.  Without supersampling the hits are the
    // pixels with alpha != 0 of the frame just read back; with it, of the sample raster (a second read-back, only when asked).
    uint64_t hits = 0;
    if (ss == 1) {
      const uint8_t* px = (const uint8_t*)h_rgba8;
      for (size_t i = 0; i < raster_bytes / 4; i++) hits += px[4 * i + 3] != 0;
    } else if (shadows) {
      std::vector<uint8_t> hi(raster_bytes);
      CLRT_CUDA(cudaMemcpyAsync(hi.data(), raster, raster_bytes, cudaMemcpyDeviceToHost, ctx->stream));
      CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
      for (size_t i = 0; i < raster_bytes / 4; i++) hits += hi[4 * i + 3] != 0;
    }
    *out_rays = (uint64_t)cam.resX * cam.resY + (shadows ? hits * (a->scene->n_lights + 1) : 0);
  }
  return 0;
}

int accel_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks) {
  // The camera travels to the kernels BY VALUE (64 bytes in the constant bank); the reference allocates and uploads a
  // CLBuffer per frame for it (BVHManager.cpp:237).
  return a->kind == ACCEL_BVH ? bvh_trace_primary(a, cam, out, rank, nranks)
       : a->kind == ACCEL_TLG ? tlg_trace_primary(a, cam, out, rank, nranks) : kd_trace_primary(a, cam, out, rank, nranks);
}
