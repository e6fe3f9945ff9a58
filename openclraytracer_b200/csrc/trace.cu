// trace.cu — BVH traversal + intersection: BVHManager::generateContacts (R/src/CLRayTracerEngine/BVHManager.cpp:234-329),
// kernels generateContacts / generateContacts2 (R/src/CLRayTracerEngine/BVHKernels.cl:84-117) ->
// bvh_generate_contact (R/Include/CLData/AccelerationStructs/BVH.h:344-408).
//
// Design (B200, no RT cores; latency/divergence bound, not HBM or tensor bound):
//  * persistent warps: the grid is sized to the resident capacity of the 148 SMs; every warp pulls 8x4-pixel packets
//    (primary) or 32-ray packets (generic) from one atomic counter -> dynamic work stealing inside the GPU, and the
//    unit the image-space multi-GPU stripes are made of (32x8-pixel tiles = 8 packets, tile t owned by rank t % nranks).
//  * 64-byte TraceNode per inner node (both child boxes + child ids, four 16-byte loads) instead of the reference's
//    48-byte node + two 32-byte child-box fetches from the children's records; 48-byte 16-byte-aligned triangles
//    instead of the per-fetch header walk of the packed scene.
//  * FAST policy: leaf children are intersected as soon as their box is accepted (so duplicate-key chains never
//    touch the stack), the nearer inner child is entered first and a child whose entry t exceeds the current best
//    is culled.  The stack only ever holds inner nodes with strictly growing key prefixes -> at most 32 entries
//    (the "short stack"), kept in local memory that stays L1 resident.
//  * REFERENCE_ORDER policy: child A before child B, no culling, leaves go through the stack — the reference's
//    exact visiting order, hence bit-identical contacts including the winners of exact ties.  Parity vehicle.
//  Box validity is the reference's in both policies: `t > 0 || origin inside` with the compare-select slab test.
#include "common.cuh"
#include "rtmath.cuh"
#include "shade_math.cuh"

namespace {

constexpr int TRACE_THREADS = 128;
constexpr int FAST_STACK = 40;
constexpr int REF_STACK = 96;
constexpr uint32_t SENTINEL = 0xFFFFFFFFu;

// `tri` is the leaf's SORTED POSITION (index into tris_sorted).  The reference's unordered A-before-B traversal tests
// leaves in ascending sorted position and accepts `0 < t < best` (BVH.h:392), so among exact ties its winner is the
// smallest position: ORDERED == false relies on visiting order (REFERENCE_ORDER policy), ORDERED == true applies the
// same rule explicitly so that a reordered traversal returns the same triangle.
struct Hit { float t; uint32_t tri; float cull; };
constexpr float CULL_SLACK = 1.00001f;   // boxes are culled against best_t * (1 + 1e-5): slab-t vs triangle-t rounding never hides a tie

template <bool ORDERED>
__device__ __forceinline__ void test_triangle(const clrt_triangle* __restrict__ tris, uint32_t id, Vec3 o, Vec3 d, Hit& h) {
  const float4* p = reinterpret_cast<const float4*>(tris + id);
  const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  const float t = triangle_intersect(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), o, d);
  if (t > 0.0f && (t < h.t || (ORDERED && t == h.t && h.tri != SENTINEL && id < h.tri))) { h.t = t; h.tri = id; h.cull = t * CULL_SLACK; }
}

struct Stats { uint32_t rays, inner, tris, lboxes; };

// ---- equal-key runs ----------------------------------------------------------------------------------------------------
// An equal-key run [start, start+count) is entered as ONE multi-triangle leaf.  The reference walks the chain node by node:
// at chain node i it tests the box of leaf i and the box of the remaining suffix, so leaf j is intersected iff its own
// box and every enclosing suffix box are valid; suffix boxes contain the leaf boxes and the slab test is monotone under
// box inclusion (also in floating point), hence "leaf box valid" alone selects the same leaves.  Boxes and triangles of
// a run are contiguous in memory (sorted order) and independent of each other - no pointer chasing.
constexpr uint32_t COOP_MIN = 6;       // runs at least this long are processed by the whole warp

template <bool STATS>
__device__ __forceinline__ void test_run_element(const clrt_triangle* __restrict__ tris, const float4* __restrict__ leafbox,
                                                 uint32_t p, Vec3 o, Vec3 d, Vec3 inv, Hit& h, Stats& st) {
  const float4 lo = __ldg(leafbox + 2 * (size_t)p), hi = __ldg(leafbox + 2 * (size_t)p + 1);
  if (STATS) st.lboxes++;
  const float tb = aabb_intersect(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, o, inv);
  const bool valid = tb > 0.0f || point_inside(lo.x, lo.y, lo.z, hi.x, hi.y, hi.z, o);
  if (valid && !(tb > h.cull)) { test_triangle<true>(tris, p, o, d, h); if (STATS) st.tris++; }
}

// Warp-cooperative run: every lane of the warp calls this; lane `leader` owns the ray and the run.  Lanes split the run's
// elements, test them against the LEADER's ray with the same arithmetic a single thread would use, and the best
// candidate is reduced with the reference's tie rule (equal t -> smaller sorted position).  Returns the merged hit in
// the leader's `h` (other lanes' `h` are untouched).
template <bool STATS>
__device__ __forceinline__ void coop_run(const clrt_triangle* __restrict__ tris, const float4* __restrict__ leafbox, int leader,
                                         uint32_t start, uint32_t count, Vec3 o, Vec3 d, Vec3 inv, Hit& h, Stats& st) {
  const int lane = threadIdx.x & 31;
  const unsigned FULL = 0xffffffffu;
  start = __shfl_sync(FULL, start, leader); count = __shfl_sync(FULL, count, leader);
  Vec3 lo_, ld_, li_;
  lo_.x = __shfl_sync(FULL, o.x, leader); lo_.y = __shfl_sync(FULL, o.y, leader); lo_.z = __shfl_sync(FULL, o.z, leader);
  ld_.x = __shfl_sync(FULL, d.x, leader); ld_.y = __shfl_sync(FULL, d.y, leader); ld_.z = __shfl_sync(FULL, d.z, leader);
  li_.x = __shfl_sync(FULL, inv.x, leader); li_.y = __shfl_sync(FULL, inv.y, leader); li_.z = __shfl_sync(FULL, inv.z, leader);
  Hit b;
  b.t = __shfl_sync(FULL, h.t, leader); b.tri = __shfl_sync(FULL, h.tri, leader); b.cull = __shfl_sync(FULL, h.cull, leader);
  for (uint32_t p = start + lane; p < start + count; p += 32) test_run_element<STATS>(tris, leafbox, p, lo_, ld_, li_, b, st);
#pragma unroll
  for (int s = 16; s > 0; s >>= 1) {
    const float ot = __shfl_xor_sync(FULL, b.t, s); const uint32_t op = __shfl_xor_sync(FULL, b.tri, s);
    if (ot < b.t || (ot == b.t && op < b.tri)) { b.t = ot; b.tri = op; }
  }
  if (lane == leader) { h.t = b.t; h.tri = b.tri; h.cull = b.t == RT_FLT_MAX ? RT_FLT_MAX : b.t * CULL_SLACK; }
}

// ---- REFERENCE_ORDER policy: the reference's visiting order, one ray per thread ---------------------------------------
template <bool STATS>
__device__ __forceinline__ Hit traverse_ref(const TraceNode* __restrict__ tnodes, const clrt_triangle* __restrict__ tris,
                                            uint32_t n, Vec3 o, Vec3 d, Stats& st) {
  Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; h.cull = RT_FLT_MAX;
  if (STATS) st.rays++;
  if (n == 0) return h;
  if (n == 1) { test_triangle<false>(tris, 0, o, d, h); if (STATS) st.tris++; return h; }    // the reference reads past its node buffer here
  const Vec3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  uint32_t stack[REF_STACK];
  int sp = 0;
  stack[sp++] = SENTINEL;
  uint32_t cur = n;                                             // root = slot N (BVHManager.cpp:254)
  do {
    if (cur >= n) {
      if (STATS) st.inner++;
      const float4* p = reinterpret_cast<const float4*>(tnodes + (cur - n));
      const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
      const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
      const float tA = aabb_intersect(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, o, inv);
      const float tB = aabb_intersect(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, o, inv);
      const bool aValid = tA > 0.0f || point_inside(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, o);
      const bool bValid = tB > 0.0f || point_inside(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, o);
      if (aValid && bValid) { cur = q3.x; if (sp < REF_STACK) stack[sp++] = q3.y; }
      else if (aValid) cur = q3.x;
      else if (bValid) cur = q3.y;
      else cur = stack[--sp];
    } else {
      test_triangle<false>(tris, cur, o, d, h);
      if (STATS) st.tris++;
      cur = stack[--sp];
    }
  } while (cur != SENTINEL);
  return h;
}

// ---- FAST policy: warp-synchronous loop, every lane of the warp must call (active == false: no ray) --------------------
// ANYHIT: returns at the first triangle with 0 < t < tmax - an occlusion query; "some hit below tmax exists" equals "the
// closest hit is below tmax", which is what the shading's shadow test asks.
// COOP: the structure holds equal-key runs of COOP_MIN or more leaves -> warp-synchronous loop with a cooperative phase;
// otherwise every lane runs its own loop (no per-step votes) and the few short runs are tested inline.
template <bool STATS, bool ANYHIT, bool COOP>
__device__ __forceinline__ Hit traverse_fast(const TraceNode* __restrict__ tnodes, const clrt_triangle* __restrict__ tris,
                                             const float4* __restrict__ leafbox, uint32_t n, bool active, Vec3 o, Vec3 d, Stats& st,
                                             float tmax = RT_FLT_MAX) {
  const unsigned FULL = 0xffffffffu;
  Hit h; h.t = tmax; h.tri = SENTINEL; h.cull = ANYHIT ? tmax * CULL_SLACK : RT_FLT_MAX;
  if (STATS && active) st.rays++;
  const Vec3 inv = v3(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
  uint32_t stack[FAST_STACK];
  int sp = 0;
  stack[sp++] = SENTINEL;
  uint32_t cur = (active && n >= 2) ? n : SENTINEL;
  if (active && n == 1) { test_triangle<false>(tris, 0, o, d, h); if (STATS) st.tris++; }
  while (COOP ? __any_sync(FULL, cur != SENTINEL) : (cur != SENTINEL)) {
    uint32_t rs0 = 0, rc0 = 0, rs1 = 0, rc1 = 0;               // runs waiting for the cooperative phase
    if (cur != SENTINEL) {
      if (STATS) st.inner++;
      const float4* p = reinterpret_cast<const float4*>(tnodes + (cur - n));
      const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
      const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
      const float tA = aabb_intersect(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, o, inv);
      const float tB = aabb_intersect(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, o, inv);
      bool aValid = (tA > 0.0f || point_inside(q0.x, q0.y, q0.z, q0.w, q1.x, q1.y, o)) && !(tA > h.cull);
      bool bValid = (tB > 0.0f || point_inside(q1.z, q1.w, q2.x, q2.y, q2.z, q2.w, o)) && !(tB > h.cull);
      const uint32_t cA = q3.x, cB = q3.y, cntA = q3.w & 0xFFFFu, cntB = q3.w >> 16;
      if (aValid && cA < n) { test_triangle<true>(tris, cA, o, d, h); if (STATS) st.tris++; aValid = false; }
      else if (aValid && cntA) {
        if (COOP && cntA >= COOP_MIN) { rs0 = q3.z; rc0 = cntA; }
        else for (uint32_t k = 0; k < cntA; k++) test_run_element<STATS>(tris, leafbox, q3.z + k, o, d, inv, h, st);
        aValid = false;
      }
      if (bValid && !(tB > h.cull) && cB < n) { test_triangle<true>(tris, cB, o, d, h); if (STATS) st.tris++; bValid = false; }
      else if (bValid && cntB) {
        if (COOP && cntB >= COOP_MIN) { if (rc0) { rs1 = cB - n; rc1 = cntB; } else { rs0 = cB - n; rc0 = cntB; } }
        else if (!(tB > h.cull)) for (uint32_t k = 0; k < cntB; k++) test_run_element<STATS>(tris, leafbox, cB - n + k, o, d, inv, h, st);
        bValid = false;
      }
      aValid = aValid && !(tA > h.cull);
      bValid = bValid && !(tB > h.cull);
      if (aValid && bValid) {
        const bool bNear = tB < tA;
        cur = bNear ? cB : cA;
        if (sp < FAST_STACK) stack[sp++] = bNear ? cA : cB;
      }
      else if (aValid) cur = cA;
      else if (bValid) cur = cB;
      else cur = stack[--sp];
    }
    // cooperative phase: the warp works off the pending runs, one owner lane at a time
    if (COOP) {
      unsigned pend = __ballot_sync(FULL, rc0 != 0);
      while (pend) {
        const int leader = __ffs(pend) - 1;
        coop_run<STATS>(tris, leafbox, leader, rs0, rc0, o, d, inv, h, st);
        if ((int)(threadIdx.x & 31) == leader) { rs0 = rs1; rc0 = rc1; rc1 = 0; }
        pend = __ballot_sync(FULL, rc0 != 0);
      }
    }
    if (ANYHIT && h.tri != SENTINEL) cur = SENTINEL;
  }
  return h;
}

__device__ __forceinline__ void write_contact(clrt_contact* __restrict__ out, uint32_t slot, uint32_t pixelIndex, const Hit& h,
                                              const clrt_triangle* __restrict__ tris) {
  uint4 head = make_uint4(pixelIndex, 0u, SENTINEL, 0u);
  float4 nd = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                 // miss: contactDist = 0 (BVH.h:402-406)
  if (h.tri != SENTINEL) {
    const float4* p = reinterpret_cast<const float4*>(tris + h.tri);
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    const Vec3 nrm = triangle_normal(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z));
    nd = make_float4(nrm.x, nrm.y, nrm.z, h.t);
    head.y = __float_as_uint(a.w);                                  // MeshHeader.materialIndex
    head.z = __float_as_uint(b.w);                                  // pad[0] = primitive (global triangle) id (Q11)
  }
  uint4* o = reinterpret_cast<uint4*>(out + slot);
  __stcs(o, head);
  __stcs(reinterpret_cast<float4*>(o) + 1, nd);
}

// statistics mode (bench / DESIGN.md figures): rays traversed, inner-node visits, triangle tests -> stats[0..2]
__device__ __forceinline__ void flush_stats(const Stats& st, unsigned long long* __restrict__ stats) {
  uint32_t r = st.rays, i = st.inner, t = st.tris, b = st.lboxes;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) {
    r += __shfl_xor_sync(0xffffffffu, r, d); i += __shfl_xor_sync(0xffffffffu, i, d); t += __shfl_xor_sync(0xffffffffu, t, d);
    b += __shfl_xor_sync(0xffffffffu, b, d);
  }
  if ((threadIdx.x & 31) == 0) {
    atomicAdd(stats, (unsigned long long)r); atomicAdd(stats + 1, (unsigned long long)i); atomicAdd(stats + 2, (unsigned long long)t);
    atomicAdd(stats + 3, (unsigned long long)b);
  }
}

// Primary rays.  Work unit: 8x4-pixel packet; 8 packets form a 32x8 tile; local tile lt of this rank is global
// tile rank + lt*nranks.  raster != 0: contact goes to out[pixelIndex] (the reference's output[c.pixelIndex] = c);
// raster == 0: stripe order, slot = lt*256 + y*32 + x, out-of-image slots get pixelIndex = UINT_MAX.
template <int MODE, bool STATS>      // MODE 0: REFERENCE_ORDER, 1: FAST, 2: FAST with cooperative runs
__global__ void __launch_bounds__(TRACE_THREADS)
trace_primary_kernel(const clrt_camera cam, const TraceNode* __restrict__ tnodes,
                     const clrt_triangle* __restrict__ tris, const float4* __restrict__ leafbox, uint32_t n, clrt_contact* __restrict__ out,
                     uint32_t rank, uint32_t nranks, int raster, uint32_t* __restrict__ counter,
                     unsigned long long* __restrict__ stats) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const uint32_t packets = stripe_tile_count(g.tiles, rank, nranks) * 8u;
  const int lane = threadIdx.x & 31;
  while (true) {
    uint32_t pk = 0;
    if (lane == 0) pk = atomicAdd(counter, 1u);
    pk = __shfl_sync(0xffffffffu, pk, 0);
    if (pk >= packets) break;
    const uint32_t lt = pk >> 3, sub = pk & 7u;
    const uint32_t gt = rank + lt * nranks;
    const uint32_t tx = (gt % g.tilesX) * CLRT_TILE_W + (sub & 3u) * 8u + (lane & 7);
    const uint32_t ty = (gt / g.tilesX) * CLRT_TILE_H + (sub >> 2) * 4u + (lane >> 3);
    const uint32_t inTile = ((sub >> 2) * 4u + (lane >> 3)) * CLRT_TILE_W + (sub & 3u) * 8u + (lane & 7);
    const bool inside = tx < cam.resX && ty < cam.resY;
    const uint32_t pixel = ty * cam.resX + tx;
    Vec3 o = v3(0.f, 0.f, 0.f), d = v3(0.f, 0.f, 1.f);
    if (inside) generate_ray(cam, pixel, o, d);
    Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; h.cull = RT_FLT_MAX;
    if (MODE == 0) { if (inside) h = traverse_ref<STATS>(tnodes, tris, n, o, d, st); }
    else h = traverse_fast<STATS, false, MODE == 2>(tnodes, tris, leafbox, n, inside, o, d, st);
    if (inside) {
      write_contact(out, raster ? pixel : lt * CLRT_TILE_PIXELS + inTile, pixel, h, tris);
    } else if (!raster) {
      Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; h.cull = RT_FLT_MAX;
      write_contact(out, lt * CLRT_TILE_PIXELS + inTile, SENTINEL, h, tris);
    }
  }
  if (STATS) flush_stats(st, stats);
}

// Generic rays (generateContacts2): contact i for ray i, pixelIndex = i.  A ray with a zero direction cannot hit
// anything in the reference either (det == 0 for every triangle); it is skipped without traversal.
template <int MODE, bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS)
trace_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const TraceNode* __restrict__ tnodes,
                  const clrt_triangle* __restrict__ tris, const float4* __restrict__ leafbox, uint32_t n, clrt_contact* __restrict__ out,
                  uint32_t* __restrict__ counter, unsigned long long* __restrict__ stats) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  const int lane = threadIdx.x & 31;
  const uint32_t packets = (count + 31u) / 32u;
  while (true) {
    uint32_t pk = 0;
    if (lane == 0) pk = atomicAdd(counter, 1u);
    pk = __shfl_sync(0xffffffffu, pk, 0);
    if (pk >= packets) break;
    const uint32_t i = pk * 32u + lane;
    float4 ro = make_float4(0.f, 0.f, 0.f, 0.f), rd = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < count) { const float4* p = reinterpret_cast<const float4*>(rays + i); ro = __ldcs(p + 1); rd = __ldcs(p + 2); }
    const bool live = i < count && !(rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f);
    Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; h.cull = RT_FLT_MAX;
    if (MODE == 0) { if (live) h = traverse_ref<STATS>(tnodes, tris, n, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), st); }
    else h = traverse_fast<STATS, false, MODE == 2>(tnodes, tris, leafbox, n, live, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), st);
    if (i < count) write_contact(out, i, i, h, tris);
  }
  if (STATS) flush_stats(st, stats);
}

// Shadow pass + shading fused (ours; the reference shades on the host and has no shadow rays): one thread per primary
// contact; for every light an occlusion query (ANYHIT) along the same shadow ray the stand-alone path builds
// (shade_math.cuh), then Blinn-Phong -> RGBA8.  Same pixels as shadow_rays_kernel + trace_rays_kernel + shade_kernel,
// without the 48-byte rays and 32-byte shadow contacts ever touching HBM.  Needs L+1 <= 32 lights.
template <bool STATS, bool COOP>
__global__ void __launch_bounds__(TRACE_THREADS)
shadow_shade_kernel(const clrt_camera cam, const clrt_contact* __restrict__ contacts, uint32_t count,
                    const clrt_light* __restrict__ lights, uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
                    const TraceNode* __restrict__ tnodes, const clrt_triangle* __restrict__ tris, const float4* __restrict__ leafbox, uint32_t n,
                    uint32_t* __restrict__ rgba8, uint32_t* __restrict__ counter, unsigned long long* __restrict__ ray_counter,
                    unsigned long long* __restrict__ stats) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  const int lane = threadIdx.x & 31;
  const uint32_t packets = (count + 31u) / 32u, L1 = n_lights + 1;
  uint32_t live_rays = 0;
  while (true) {
    uint32_t pk = 0;
    if (lane == 0) pk = atomicAdd(counter, 1u);
    pk = __shfl_sync(0xffffffffu, pk, 0);
    if (pk >= packets) break;
    const uint32_t i = pk * 32u + lane;
    uint4 head = make_uint4(SENTINEL, 0u, SENTINEL, 0u);
    float4 nd = make_float4(0.f, 0.f, 0.f, 0.f);
    if (i < count) {
      head = __ldcs(reinterpret_cast<const uint4*>(contacts + i));
      nd = __ldcs(reinterpret_cast<const float4*>(contacts + i) + 1);
    }
    const bool live = nd.w > 0.0f && head.x != SENTINEL;
    uint32_t occl = 0;
    Vec3 o = v3(0.f, 0.f, 0.f), P = v3(0.f, 0.f, 0.f);
    if (live) { Vec3 d; generate_ray(cam, head.x, o, d); P = vadd(o, vscale(d, nd.w)); live_rays += L1; }
    for (uint32_t l = 0; l < L1; l++) {                       // warp-uniform: every lane takes part in the traversal loop
      Vec3 org = v3(0.f, 0.f, 0.f), dir = v3(0.f, 0.f, 1.f);
      float limit = 0.0f;
      if (live) {
        float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
        limit = vlength(vsub(lp, P)) - SHADOW_EPS_REL * nd.w;
        shadow_ray(cam, head.x, nd.w, lights, n_lights, l, org, dir);
      }
      const Hit h = traverse_fast<STATS, true, COOP>(tnodes, tris, leafbox, n, live, org, dir, st, limit);
      occl |= (h.tri != SENTINEL ? 1u : 0u) << l;
    }
    if (i < count) {
      uint32_t packed;
      shade_contact(cam, head, nd, lights, n_lights, materials, n_materials, [&](uint32_t l, float) { return ((occl >> l) & 1u) != 0u; }, packed);
      __stcs(rgba8 + i, packed);
    }
  }
  if (ray_counter) {
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) live_rays += __shfl_xor_sync(0xffffffffu, live_rays, dd);
    if (lane == 0 && live_rays) atomicAdd(ray_counter, (unsigned long long)live_rays);
  }
  if (STATS) flush_stats(st, stats);
}

template <typename K>
int persistent_grid(clrt_ctx* ctx, K kernel, int threads, unsigned* grid) {
  int per_sm = 0;
  CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, 0));
  if (per_sm < 1) per_sm = 1;
  *grid = (unsigned)(ctx->sm_count * per_sm);
  return 0;
}

} // namespace

template <int M, bool S>
static int launch_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks, int raster, uint32_t packets) {
  unsigned grid = 0;
  CLRT_TRY(persistent_grid(a->ctx, trace_primary_kernel<M, S>, TRACE_THREADS, &grid));
  grid = min(grid, (packets + 3) / 4);
  trace_primary_kernel<M, S><<<grid, TRACE_THREADS, 0, a->ctx->stream>>>(cam, a->d_tnodes, a->d_tris_sorted, a->d_leafbox_sorted, a->n, out, rank, nranks, raster,
                                                                        a->d_tile_counter, (unsigned long long*)(a->d_tile_counter + 4));
  CLRT_CHECK_LAUNCH();
  return 0;
}
template <int M, bool S>
static int launch_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count, uint32_t packets) {
  unsigned grid = 0;
  CLRT_TRY(persistent_grid(a->ctx, trace_rays_kernel<M, S>, TRACE_THREADS, &grid));
  grid = min(grid, (packets + 3) / 4);
  trace_rays_kernel<M, S><<<grid, TRACE_THREADS, 0, a->ctx->stream>>>(rays, count, a->d_tnodes, a->d_tris_sorted, a->d_leafbox_sorted, a->n, out,
                                                                     a->d_tile_counter, (unsigned long long*)(a->d_tile_counter + 4));
  CLRT_CHECK_LAUNCH();
  return 0;
}
// 0: REFERENCE_ORDER, 1: FAST, 2: FAST + cooperative runs (the structure holds runs of COOP_MIN or more equal keys)
static int trace_mode_index(const clrt_accel* a) {
  if (a->trace_mode == CLRT_TRACE_REFERENCE_ORDER) return 0;
  return a->max_run >= COOP_MIN ? 2 : 1;
}

int bvh_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks) {
  const uint32_t resX = cam.resX, resY = cam.resY;
  clrt_ctx* ctx = a->ctx;
  if (!a->built) CLRT_FAIL("generateContacts: the BVH was not constructed");
  const bool raster = (nranks == 0);                 // nranks == 0: whole frame, raster-ordered output
  if (raster) { rank = 0; nranks = 1; }
  const TileGrid g = make_tile_grid(resX, resY);
  const uint32_t packets = stripe_tile_count(g.tiles, rank, nranks) * 8u;
  if (!packets) return 0;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  const int m = trace_mode_index(a);
  if (a->stats) return m == 0 ? launch_primary<0, true>(a, cam, out, rank, nranks, raster, packets)
                     : m == 1 ? launch_primary<1, true>(a, cam, out, rank, nranks, raster, packets)
                              : launch_primary<2, true>(a, cam, out, rank, nranks, raster, packets);
  return m == 0 ? launch_primary<0, false>(a, cam, out, rank, nranks, raster, packets)
       : m == 1 ? launch_primary<1, false>(a, cam, out, rank, nranks, raster, packets)
                : launch_primary<2, false>(a, cam, out, rank, nranks, raster, packets);
}

int bvh_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count) {
  clrt_ctx* ctx = a->ctx;
  if (!a->built) CLRT_FAIL("generateContacts: the BVH was not constructed");
  if (!count) return 0;
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  const int m = trace_mode_index(a);
  if (a->stats) return m == 0 ? launch_rays<0, true>(a, rays, out, count, packets) : m == 1 ? launch_rays<1, true>(a, rays, out, count, packets)
                                                                                            : launch_rays<2, true>(a, rays, out, count, packets);
  return m == 0 ? launch_rays<0, false>(a, rays, out, count, packets) : m == 1 ? launch_rays<1, false>(a, rays, out, count, packets)
                                                                               : launch_rays<2, false>(a, rays, out, count, packets);
}

template <bool S, bool C>
static int launch_shadow_shade(clrt_accel* a, const clrt_camera& cam, const clrt_contact* contacts, uint32_t count, uint32_t* rgba8,
                               unsigned long long* d_ray_counter, uint32_t packets) {
  const clrt_scene* sc = a->scene;
  unsigned grid = 0;
  CLRT_TRY(persistent_grid(a->ctx, shadow_shade_kernel<S, C>, TRACE_THREADS, &grid));
  grid = min(grid, (packets + 3) / 4);
  shadow_shade_kernel<S, C><<<grid, TRACE_THREADS, 0, a->ctx->stream>>>(cam, contacts, count, sc->d_lights, sc->n_lights, sc->d_materials, sc->n_materials,
      a->d_tnodes, a->d_tris_sorted, a->d_leafbox_sorted, a->n, rgba8, a->d_tile_counter, d_ray_counter, (unsigned long long*)(a->d_tile_counter + 4));
  CLRT_CHECK_LAUNCH();
  return 0;
}

// Fused shadow + shade over `count` contacts (BVH only).  d_ray_counter (may be NULL) accumulates the live shadow rays.
int bvh_shadow_shade(clrt_accel* a, const clrt_camera& cam, const clrt_contact* contacts, uint32_t count, uint32_t* rgba8,
                     unsigned long long* d_ray_counter) {
  clrt_ctx* ctx = a->ctx;
  if (!a->built) CLRT_FAIL("shadow pass: the BVH was not constructed");
  if (!count) return 0;
  if (a->scene->n_lights + 1 > 32) CLRT_FAIL("fused shadow pass supports at most 31 scene lights");
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  const bool coop = a->max_run >= COOP_MIN;
  if (a->stats) return coop ? launch_shadow_shade<true, true>(a, cam, contacts, count, rgba8, d_ray_counter, packets)
                            : launch_shadow_shade<true, false>(a, cam, contacts, count, rgba8, d_ray_counter, packets);
  return coop ? launch_shadow_shade<false, true>(a, cam, contacts, count, rgba8, d_ray_counter, packets)
              : launch_shadow_shade<false, false>(a, cam, contacts, count, rgba8, d_ray_counter, packets);
}
