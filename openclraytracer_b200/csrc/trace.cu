// trace.cu — BVH traversal + intersection: BVHManager::generateContacts (R/src/CLRayTracerEngine/BVHManager.cpp:234-329),
// kernels generateContacts / generateContacts2 (R/src/CLRayTracerEngine/BVHKernels.cl:84-117) ->
// bvh_generate_contact (R/Include/CLData/AccelerationStructs/BVH.h:344-408).
//
// Design (B200, no RT cores; latency/divergence bound, not HBM or tensor bound):
//  * persistent warps: the grid is sized to the resident capacity of the 148 SMs; every warp pulls 8x4-pixel packets
//    (primary) or 32-ray packets (generic) from one atomic counter -> dynamic work stealing inside the GPU, and the
//    unit the image-space multi-GPU stripes are made of (32x8-pixel tiles = 8 packets, tile t owned by rank t % nranks).
//  * 48-byte 16-byte-aligned triangles in tree order instead of the per-fetch header walk of the packed scene.
//  * FAST policy: 128-byte 4-wide nodes (WideNode: four child boxes SoA + four child ids, eight 16-byte loads) — the
//    2-level collapse of the binary tree, so a ray's dependent load chain is half as long; leaf children are intersected
//    as soon as their box is accepted, inner children are entered nearest first, the others wait on a stack of
//    (entry distance, node) pairs in local memory and are culled against the current best when popped.
//  * The tree FAST walks is the reference tree, or the separate trace tree (trace_tree.cu) when the reference tree has
//    long equal-key chains or collapsed Morton axes.  Exact ties go to the reference's winner in both.
//  * REFERENCE_ORDER policy: the binary reference topology through 64-byte TraceNodes (both child boxes + child ids in
//    one record), child A before child B, no culling, leaves go through the stack — the reference's exact visiting order, hence bit-identical contacts including the winners of exact ties.  Parity vehicle.
//  Box validity is the reference's in both policies: `t > 0 || origin inside` with the compare-select slab test.
#include <stdlib.h>
#include "common.cuh"
#include "rtmath.cuh"
#include "shade_math.cuh"

namespace {

#ifndef CLRT_TRACE_THREADS
#define CLRT_TRACE_THREADS 128
#endif
constexpr int TRACE_THREADS = CLRT_TRACE_THREADS;
constexpr int FAST_STACK = 96;     // (entry distance, node) pairs; a visit pushes at most three: 3 x wide depth, and a trace tree of 2^30 leaves is at most 60 binary = 30 wide levels deep
constexpr int REF_STACK = 96;
constexpr uint32_t SENTINEL = 0xFFFFFFFFu;

// `tri` is the leaf's SORTED POSITION (index into tris_sorted).  The reference's unordered A-before-B traversal tests
// leaves in ascending sorted position and accepts `0 < t < best` (BVH.h:392), so among exact ties its winner is the
// smallest position: ORDERED == false relies on visiting order (REFERENCE_ORDER policy), ORDERED == true applies the
// same rule explicitly so that a reordered traversal returns the same triangle.
struct Hit { float t; uint32_t tri; float cull; uint32_t ref; };
constexpr float CULL_SLACK = 1.00001f;   // boxes are culled against best_t * (1 + 1e-5): slab-t vs triangle-t rounding never hides a tie

template <bool ORDERED>
__device__ __forceinline__ void test_triangle(const clrt_triangle* __restrict__ tris, uint32_t id, Vec3 o, Vec3 d, Hit& h) {
  const float4* p = reinterpret_cast<const float4*>(tris + id);
  const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
  const float t = triangle_intersect(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), o, d);
  // c.w = the leaf's position in the REFERENCE's sorted order (== id when the traversal tree is the reference tree)
  if (t > 0.0f && (t < h.t || (ORDERED && t == h.t && h.tri != SENTINEL && __float_as_uint(c.w) < h.ref))) {
    h.t = t; h.tri = id; h.cull = t * CULL_SLACK; h.ref = __float_as_uint(c.w);
  }
}

struct Stats { uint32_t rays, inner, tris, lboxes; };

// ---- REFERENCE_ORDER policy: the reference's visiting order, one ray per thread ---------------------------------------
template <bool STATS>
__device__ __forceinline__ Hit traverse_ref(const TraceNode* __restrict__ tnodes, const clrt_triangle* __restrict__ tris,
                                            uint32_t n, Vec3 o, Vec3 d, Stats& st, uint32_t* __restrict__ overflow) {
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
      if (aValid && bValid) { cur = q3.x; if (sp < REF_STACK) stack[sp++] = q3.y; else atomicAdd(overflow, 1u); }   // counted: clrt_accel_check fails the call
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

// ---- FAST policy ---------------------------------------------------------------------------------------------------------
// ANYHIT template argument: 0 closest hit, 1 occlusion query, 2 decided per ray (Lane::any — the fused frame kernel traces both kinds).
#define LANE_ANY(L) (ANYHIT == 1 || (ANYHIT == 2 && (L).any))
// Per-lane traversal state.  ANYHIT: an occlusion query - ends at the first triangle with 0 < t < tmax; "some hit below
// tmax exists" equals "the closest hit is below tmax", which is what the shading's shadow test asks.
struct Lane {
  Vec3 o, d;
#if CLRT_WIDE_Q16
  Vec3 An, Af, B;                 // slab distances of grid position q: near bounds q * B + An, far bounds q * B + Af (lane_begin)
#else
  Vec3 inv;
#endif
  Hit h;
  uint32_t cur;
  int sp;
  uint32_t visits;                // statistics mode only: inner-node visits of the current ray (-> Contact.pad[1])
  uint32_t q0, q1, q2, q3, qn;    // postponed leaves (positions in the triangle array) and their count
  bool any;                       // ANYHIT == 2 only: this ray is an occlusion query
};

// Compile-time shape of the traversal kernels (profiles/variants.sh builds and measures the alternatives):
//   CLRT_TRACE_THREADS  threads per CTA
//   CLRT_SSTACK         traversal-stack entries per thread kept in SHARED memory (entry-major: entry e of thread t at
//                       [e * threads + t], 8-byte words -> conflict free); deeper entries spill to a local-memory array that a
//                       balanced tree never reaches (Soup-1M: deepest stack 19)
//   CLRT_TOP_STAGED     records of the top table (common.cuh: the first levels of the wide tree in breadth-first order) copied
//                       into shared memory by ONE bulk-async (TMA) copy per CTA at kernel start: 21 = 3 levels (2.6 KB),
//                       85 = 4 levels (10.6 KB), 341 = 5 levels (42.6 KB); 0 = the table is read from global memory
#ifndef CLRT_TRACE_THREADS
#define CLRT_TRACE_THREADS 128
#endif
#ifndef CLRT_SSTACK
#define CLRT_SSTACK 16
#endif
#ifndef CLRT_TOP_STAGED
#define CLRT_TOP_STAGED 85
#endif
#ifndef CLRT_UNIFORM_STACK
#define CLRT_UNIFORM_STACK 0
#endif
#ifndef CLRT_MIN_CTAS
#define CLRT_MIN_CTAS 6          // __launch_bounds__ second argument of the FAST kernels (register cap = 65536 / (threads x CTAs))
#endif
#ifndef CLRT_PREFETCH_TOP
#define CLRT_PREFETCH_TOP 0      // prefetch.global.L1 of the node on top of the stack (the one visited after the nearest child's subtree)
#endif
constexpr int SSTACK = CLRT_SSTACK;
constexpr int TOP_STAGED = CLRT_TOP_STAGED;
static_assert(TRACE_THREADS == CLRT_TRACE_THREADS, "TRACE_THREADS");
static_assert(TOP_STAGED == 0 || TOP_STAGED == 5 || TOP_STAGED == 21 || TOP_STAGED == 85 || TOP_STAGED == 341, "CLRT_TOP_STAGED: whole levels");
static_assert(SSTACK >= 4 && SSTACK <= FAST_STACK, "CLRT_SSTACK");
constexpr size_t FAST_SMEM_BYTES = (size_t)TOP_STAGED * sizeof(WideNode) + (size_t)SSTACK * TRACE_THREADS * sizeof(uint2) + 16;

struct StackRef {
  uint2* s;        // shared-memory column of this thread (stride TRACE_THREADS)
  uint2* l;        // local-memory spill [FAST_STACK - SSTACK]
  uint32_t* overflow;
};

template <bool STATS, int ANYHIT>
__device__ __forceinline__ void lane_begin(Lane& L, const clrt_triangle* __restrict__ tris, uint32_t n, const QFrame* __restrict__ qf,
                                           Vec3 o, Vec3 d, float tmax, Stats& st) {
  L.o = o; L.d = d;
  // 1/d as the reference computes it, except that an exactly zero component gives +-1e30 instead of +-inf: (bound - origin) * inv
  // is then 0 when the origin lies ON the bound (the reference's 0 * inf = NaN is dropped by its compare-select min/max,
  // AABB.h:126-142, i.e. the slab is accepted — same outcome) and +-huge otherwise (rejected like +-inf).  No NaN reaches the
  // fminf/fmaxf of the box test, which would turn it into a rejection of both boxes along a seam in the ray's plane.
  const float big = 1.0e30f;
  const Vec3 inv = v3(d.x == 0.0f ? copysignf(big, d.x) : 1.0f / d.x, d.y == 0.0f ? copysignf(big, d.y) : 1.0f / d.y, d.z == 0.0f ? copysignf(big, d.z) : 1.0f / d.z);
#if CLRT_WIDE_Q16
  // Grid position q (16 bit) on an axis lies at lo + q * step, i.e. at ray distance A + q * B with A = (lo - origin) * inv and
  // B = step * inv.  q reaches the FMA as the float 2^23 + q (one PRMT builds its bits, no conversion), so the constant is
  // A0 = A - 2^23 * B.  Rounding: A0 is rounded at magnitude 2^23 |B| + |A| (half a step of error), q * B exactly cancels the
  // 2^23 * B part because the same rounded B is used, A carries 2^-23 |A|.  Near bounds are therefore moved 0.55 |B| + 2^-21 |A|
  // earlier and far bounds as much later: the tested box still contains the stored one (which contains the exact one).
  if (n >= 2) {
    const float4 f0 = __ldg(reinterpret_cast<const float4*>(qf)), f1 = __ldg(reinterpret_cast<const float4*>(qf) + 1);
    const Vec3 A = v3((f0.x - o.x) * inv.x, (f0.y - o.y) * inv.y, (f0.z - o.z) * inv.z);
    L.B = v3(f1.x * inv.x, f1.y * inv.y, f1.z * inv.z);
    const Vec3 A0 = v3(__fmaf_rn(-8388608.0f, L.B.x, A.x), __fmaf_rn(-8388608.0f, L.B.y, A.y), __fmaf_rn(-8388608.0f, L.B.z, A.z));
    const float k = 1.0f / 2097152.0f;
    const Vec3 w = v3(0.55f * fabsf(L.B.x) + k * fabsf(A.x), 0.55f * fabsf(L.B.y) + k * fabsf(A.y), 0.55f * fabsf(L.B.z) + k * fabsf(A.z));
    L.An = v3(A0.x - w.x, A0.y - w.y, A0.z - w.z);
    L.Af = v3(A0.x + w.x, A0.y + w.y, A0.z + w.z);
  }
#else
  L.inv = inv;
#endif
  L.h.t = tmax; L.h.tri = SENTINEL; L.h.cull = LANE_ANY(L) ? tmax * CULL_SLACK : RT_FLT_MAX;
  L.sp = 0; L.qn = 0;
  L.cur = n >= 2 ? WIDE_TOP_BIT : SENTINEL;                  // slot 0 of the top table = the root
  if (STATS) { st.rays++; L.visits = 0; }
  if (n == 1) { test_triangle<false>(tris, 0, o, d, L.h); if (STATS) st.tris++; }     // the reference reads past its node buffer here
}

// One visit of a 4-wide node (WideNode, common.cuh: the 2-level collapse of the binary trace structure written by
// widen_kernel, trace_tree.cu): ONE dependent 128-byte fetch decides four boxes, which halves the length of the dependent
// load chain a ray walks.  Accepted leaf children go to the lane's leaf queue (tested later, warp-wide: flush_one); valid
// inner children are sorted by entry distance, the nearest is entered, the others go to the stack WITH their entry
// distance so that a pop can cull them against the hit found meanwhile.  Box validity is the reference's
// (child_entry below); children beyond best_t * (1 + 1e-5) are culled.
#define T_INVALID __int_as_float(0x7f800000)      /* +inf marks an empty / rejected slot */

// Entry distance of one child box, or T_INVALID.  Same six products (bound - origin) * (1/d) as the reference's AABBIntersect
// (AABB.h:126-142), so for finite 1/d tn / tf are the reference's tmin / tmax bit for bit; the per-axis near/far pick is a
// min/max instead of the sign select, the three pairwise overlap flags are `tn <= tf`, and the reference's validity
// `t > 0 || origin inside` (AABB.h:178-184) is `tf >= 0` (origin inside <=> every axis has near <= 0 <= far).  Axis-parallel
// rays: see lane_begin.  Empty slots hold NaN boxes (widen_kernel): every comparison fails.
#if !CLRT_WIDE_Q16
__device__ __forceinline__ float child_entry(float x0, float y0, float z0, float x1, float y1, float z1, Vec3 o, Vec3 inv, float cull) {
  const float ax = (x0 - o.x) * inv.x, bx = (x1 - o.x) * inv.x;
  const float ay = (y0 - o.y) * inv.y, by = (y1 - o.y) * inv.y;
  const float az = (z0 - o.z) * inv.z, bz = (z1 - o.z) * inv.z;
  const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz));
  const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  return (tn <= fminf(tf, cull) && tf >= 0.0f) ? tn : T_INVALID;
}
#endif
__device__ __forceinline__ void cswap(float& ta, uint32_t& ca, float& tb, uint32_t& cb) {
  if (tb < ta) { const float t = ta; ta = tb; tb = t; const uint32_t c = ca; ca = cb; cb = c; }
}

// The stack of (entry distance, node): the first SSTACK entries in shared memory, the rest in local memory.  A push beyond
// FAST_STACK cannot be honoured; it is COUNTED (clrt_accel_check fails the call) instead of silently dropping a subtree.
__device__ __forceinline__ void stack_push(Lane& L, const StackRef& S, float t, uint32_t c) {
  const uint2 e = make_uint2(__float_as_uint(t), c);
  if (L.sp < SSTACK) S.s[L.sp * TRACE_THREADS] = e;
  else if (L.sp < FAST_STACK) S.l[L.sp - SSTACK] = e;
  else { atomicAdd(S.overflow, 1u); return; }
  L.sp++;
}
__device__ __forceinline__ uint2 stack_pop(Lane& L, const StackRef& S) {      // L.sp > 0
  --L.sp;
  return L.sp < SSTACK ? S.s[L.sp * TRACE_THREADS] : S.l[L.sp - SSTACK];
}

template <bool STATS, int ANYHIT>
__device__ __forceinline__ void lane_visit(Lane& L, const WideNode* __restrict__ wnodes, const WideNode* s_top, const clrt_triangle* __restrict__ tris,
                                           uint32_t n, Stats& st, const StackRef& S) {
  if (STATS) { st.inner++; L.visits++; }
  const Vec3 o = L.o, d = L.d;
  // top-table slots below TOP_STAGED are read from the CTA's shared-memory copy, everything else from global memory
  const uint32_t slot = L.cur & ~WIDE_TOP_BIT;
  const WideNode* node = (L.cur & WIDE_TOP_BIT) ? ((TOP_STAGED && slot < (uint32_t)TOP_STAGED) ? s_top + slot : wnodes + n + slot)
                                                : wnodes + (L.cur - n);
#if CLRT_WIDE_Q16
  const uint4* p = reinterpret_cast<const uint4*>(node);
  const uint4 qa = p[0], qb = p[1], qc = p[2], ch = p[3];
  uint32_t c0 = ch.x, c1 = ch.y, c2 = ch.z, c3 = ch.w;
  float t0, t1, t2, t3;
  {
    // near / far grid positions per axis by the sign of the direction (B carries the sign of 1/d)
    const bool nx = L.B.x < 0.0f, ny = L.B.y < 0.0f, nz = L.B.z < 0.0f;
    const uint32_t nxa = nx ? qb.z : qa.x, nxb = nx ? qb.w : qa.y, fxa = nx ? qa.x : qb.z, fxb = nx ? qa.y : qb.w;
    const uint32_t nya = ny ? qc.x : qa.z, nyb = ny ? qc.y : qa.w, fya = ny ? qa.z : qc.x, fyb = ny ? qa.w : qc.y;
    const uint32_t nza = nz ? qc.z : qb.x, nzb = nz ? qc.w : qb.y, fza = nz ? qb.x : qc.z, fzb = nz ? qb.y : qc.w;
    const float cull = L.h.cull;
    auto entry = [&](uint32_t wnx, uint32_t wny, uint32_t wnz, uint32_t wfx, uint32_t wfy, uint32_t wfz, unsigned sel) {
      // 0x4B000000 | q is the float 2^23 + q: bytes (q.lo, q.hi, 0x00, 0x4B) in one PRMT
      const float tn = fmaxf(fmaxf(__fmaf_rn(__uint_as_float(__byte_perm(wnx, 0x4B000000u, sel)), L.B.x, L.An.x),
                                   __fmaf_rn(__uint_as_float(__byte_perm(wny, 0x4B000000u, sel)), L.B.y, L.An.y)),
                             __fmaf_rn(__uint_as_float(__byte_perm(wnz, 0x4B000000u, sel)), L.B.z, L.An.z));
      const float tf = fminf(fminf(__fmaf_rn(__uint_as_float(__byte_perm(wfx, 0x4B000000u, sel)), L.B.x, L.Af.x),
                                   __fmaf_rn(__uint_as_float(__byte_perm(wfy, 0x4B000000u, sel)), L.B.y, L.Af.y)),
                             __fmaf_rn(__uint_as_float(__byte_perm(wfz, 0x4B000000u, sel)), L.B.z, L.Af.z));
      return (tn <= fminf(tf, cull) && tf >= 0.0f) ? tn : T_INVALID;
    };
    t0 = entry(nxa, nya, nza, fxa, fya, fza, 0x7410u);
    t1 = entry(nxa, nya, nza, fxa, fya, fza, 0x7432u);
    t2 = entry(nxb, nyb, nzb, fxb, fyb, fzb, 0x7410u);
    t3 = entry(nxb, nyb, nzb, fxb, fyb, fzb, 0x7432u);
    // an empty slot holds an inverted box, which the test rejects for every finite ray; its id says so too
    if (c0 == SENTINEL) t0 = T_INVALID;
    if (c1 == SENTINEL) t1 = T_INVALID;
    if (c2 == SENTINEL) t2 = T_INVALID;
    if (c3 == SENTINEL) t3 = T_INVALID;
  }
#else
  const Vec3 inv = L.inv;
  const float4* p = reinterpret_cast<const float4*>(node);
  const float4 lx = p[0], ly = p[1], lz = p[2], hx = p[3], hy = p[4], hz = p[5];
  const uint4 ch = *reinterpret_cast<const uint4*>(p + 6);
  uint32_t c0 = ch.x, c1 = ch.y, c2 = ch.z, c3 = ch.w;
  float t0 = child_entry(lx.x, ly.x, lz.x, hx.x, hy.x, hz.x, o, inv, L.h.cull);
  float t1 = child_entry(lx.y, ly.y, lz.y, hx.y, hy.y, hz.y, o, inv, L.h.cull);
  float t2 = child_entry(lx.z, ly.z, lz.z, hx.z, hy.z, hz.z, o, inv, L.h.cull);
  float t3 = child_entry(lx.w, ly.w, lz.w, hx.w, hy.w, hz.w, o, inv, L.h.cull);
#endif
  // Leaves are POSTPONED: their ids go to a four-entry queue in registers and the triangle tests run warp-wide in flush_one once
  // enough lanes hold one.  Tested on the spot, the Moeller-Trumbore code ran with 2.7 of 32 lanes (round-1 ncu source view:
  // only ~9 % of the lane-steps find a leaf, but with 16+ active lanes almost every warp-step found one somewhere).  A fifth
  // leaf is tested at once.
  {
    uint32_t leaves = (c0 < n && t0 != T_INVALID ? 1u : 0u) | (c1 < n && t1 != T_INVALID ? 2u : 0u) |
                      (c2 < n && t2 != T_INVALID ? 4u : 0u) | (c3 < n && t3 != T_INVALID ? 8u : 0u);
    if (leaves & 1u) t0 = T_INVALID;
    if (leaves & 2u) t1 = T_INVALID;
    if (leaves & 4u) t2 = T_INVALID;
    if (leaves & 8u) t3 = T_INVALID;
    while (leaves) {
      const uint32_t id = (leaves & 1u) ? c0 : (leaves & 2u) ? c1 : (leaves & 4u) ? c2 : c3;
      leaves &= leaves - 1u;
      if (L.qn < 4u) { L.q3 = L.q2; L.q2 = L.q1; L.q1 = L.q0; L.q0 = id; L.qn++; }
      else { test_triangle<true>(tris, id, o, d, L.h); if (STATS) st.tris++; }
    }
  }
  if (LANE_ANY(L) && L.h.tri != SENTINEL) { L.cur = SENTINEL; L.sp = 0; L.qn = 0; return; }
  if (t0 > L.h.cull) t0 = T_INVALID;
  if (t1 > L.h.cull) t1 = T_INVALID;
  if (t2 > L.h.cull) t2 = T_INVALID;
  if (t3 > L.h.cull) t3 = T_INVALID;
  cswap(t0, c0, t1, c1); cswap(t2, c2, t3, c3); cswap(t0, c0, t2, c2); cswap(t1, c1, t3, c3); cswap(t1, c1, t2, c2);   // ascending entry distance
#if CLRT_UNIFORM_STACK
  // Branch-free: every valid child is pushed, farthest first (an invalid one is stored too, but the stack pointer does not move),
  // and the next node is ALWAYS popped: the nearest child when there is one (its entry distance was just accepted, so the cull test
  // below passes), else older entries until one survives the cull distance.  One shared-memory round trip on the critical path
  // against four divergent branches: the push / pop code ran with 5-6 of 32 lanes (profiles/r02a_source_hotspots.md).
  if (L.sp + 4 <= SSTACK) {
    uint2* top = S.s + L.sp * TRACE_THREADS;
    top[0] = make_uint2(__float_as_uint(t3), c3); top += (t3 != T_INVALID) ? TRACE_THREADS : 0; L.sp += (t3 != T_INVALID);
    top[0] = make_uint2(__float_as_uint(t2), c2); top += (t2 != T_INVALID) ? TRACE_THREADS : 0; L.sp += (t2 != T_INVALID);
    top[0] = make_uint2(__float_as_uint(t1), c1); top += (t1 != T_INVALID) ? TRACE_THREADS : 0; L.sp += (t1 != T_INVALID);
    top[0] = make_uint2(__float_as_uint(t0), c0); L.sp += (t0 != T_INVALID);
  } else {
    if (t3 != T_INVALID) stack_push(L, S, t3, c3);
    if (t2 != T_INVALID) stack_push(L, S, t2, c2);
    if (t1 != T_INVALID) stack_push(L, S, t1, c1);
    if (t0 != T_INVALID) stack_push(L, S, t0, c0);
  }
#else
  if (t3 != T_INVALID) stack_push(L, S, t3, c3);
  if (t2 != T_INVALID) stack_push(L, S, t2, c2);
  if (t1 != T_INVALID) stack_push(L, S, t1, c1);
#if CLRT_PREFETCH_TOP
  if (t1 != T_INVALID && !(c1 & WIDE_TOP_BIT)) asm volatile("prefetch.global.L1 [%0];" :: "l"(wnodes + (c1 - n)));
#endif
  if (t0 != T_INVALID) { L.cur = c0; return; }
#endif
  L.cur = SENTINEL;
  while (L.sp > 0) {
    const uint2 e = stack_pop(L, S);
    if (!(__uint_as_float(e.x) > L.h.cull)) { L.cur = e.y; break; }
  }
}

// One postponed triangle test per lane that holds one (the queue is LIFO: q0 is the newest).  A closest-hit ray goes on with the
// shorter cull distance; an occlusion query that hit is complete.
template <bool STATS, int ANYHIT>
__device__ __forceinline__ void flush_one(Lane& L, const clrt_triangle* __restrict__ tris, Stats& st) {
  if (L.qn != 0u) {
    const uint32_t id = L.q0;
    L.q0 = L.q1; L.q1 = L.q2; L.q2 = L.q3; L.qn--;
    test_triangle<true>(tris, id, L.o, L.d, L.h);
    if (STATS) st.tris++;
    if (LANE_ANY(L) && L.h.tri != SENTINEL) { L.qn = 0u; L.cur = SENTINEL; L.sp = 0; }
  }
}

// Persistent-warp driver with lane refill.  A "unit" is a pixel slot / ray / contact; units come from one global atomic
// counter in warp-private batches of BATCH consecutive indices (work stealing between all warps of the GPU).  Every step
// has two phases, each run by all lanes that have something to do in it:
//   A  lane_visit: lanes with a current node fetch it, test its four boxes, queue accepted leaves, push / pop;
//   B  flush_one:  once FLUSH_MIN lanes hold a postponed leaf (or a queue is QFULL deep, or no lane can do phase A) every
//      lane with a queue tests ONE triangle.
// A lane whose ray is complete waits; when REFILL_MIN lanes wait, they complete their units together (Job::finish may aim
// the unit's next ray) and idle lanes are handed the next units of the warp's batch, so lanes never wait for the slowest ray
// of a packet (rays of one warp differ several-fold in node visits).
//   Job::start(unit, L)  -> true: a ray is ready in L;  false: the unit completed without tracing (output written)
//   Job::finish(L)       -> true: the unit continues with another ray (L re-initialised); false: unit complete
#ifndef CLRT_WINDOW
#define CLRT_WINDOW 2
#endif
#ifndef CLRT_REFILL_MIN
#define CLRT_REFILL_MIN 12
#endif
#ifndef CLRT_FLUSH_MIN
#define CLRT_FLUSH_MIN 8
#endif
#ifndef CLRT_QFULL
#define CLRT_QFULL 3
#endif
#ifndef CLRT_BATCH
#define CLRT_BATCH 32
#endif
constexpr int WINDOW = CLRT_WINDOW;
constexpr int REFILL_MIN = CLRT_REFILL_MIN;
constexpr int FLUSH_MIN = CLRT_FLUSH_MIN;
constexpr uint32_t QFULL = CLRT_QFULL;
constexpr uint32_t BATCH = CLRT_BATCH;

// Shared memory of a FAST kernel: [top table prefix | stack columns | mbarrier].  One thread arms the barrier and issues the
// bulk copy (cp.async.bulk = the TMA engine's linear mode); everybody waits on the barrier's phase 0.
__device__ __forceinline__ const WideNode* stage_top_table(unsigned char* smem, const WideNode* __restrict__ g_top, uint32_t n) {
  WideNode* s_top = reinterpret_cast<WideNode*>(smem);
  if (TOP_STAGED && n >= 2) {                                  // n < 2: no wide nodes exist (lane_begin answers such scenes directly)
    unsigned long long* bar = reinterpret_cast<unsigned long long*>(smem + (size_t)TOP_STAGED * sizeof(WideNode) + (size_t)SSTACK * TRACE_THREADS * sizeof(uint2));
    const uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(bar), dst_a = (uint32_t)__cvta_generic_to_shared(s_top);
    constexpr uint32_t bytes = (uint32_t)(TOP_STAGED * sizeof(WideNode));
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar_a) : "memory");
      asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar_a), "r"(bytes) : "memory");
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                   :: "r"(dst_a), "l"(g_top), "r"(bytes), "r"(bar_a) : "memory");
    }
    uint32_t done = 0;
    while (!done)
      asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], 0; selp.u32 %0, 1, 0, p; }" : "=r"(done) : "r"(bar_a) : "memory");
  }
  return s_top;
}

template <class Job, bool STATS, int ANYHIT>
__device__ __forceinline__ void run_jobs(Job& job, const WideNode* __restrict__ wnodes, const clrt_triangle* __restrict__ tris,
                                         uint32_t n, uint32_t total, uint32_t* __restrict__ counter, Stats& st, uint32_t* __restrict__ overflow) {
  extern __shared__ __align__(128) unsigned char fast_smem[];
  const WideNode* s_top = stage_top_table(fast_smem, wnodes + n, n);
  uint2 spill[FAST_STACK - SSTACK > 0 ? FAST_STACK - SSTACK : 1];
  const StackRef S{reinterpret_cast<uint2*>(fast_smem + (size_t)TOP_STAGED * sizeof(WideNode)) + threadIdx.x, spill, overflow};
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  Lane L; L.cur = SENTINEL; L.sp = 0; L.qn = 0;
  bool busy = false, exhausted = false;
  uint32_t pool_next = 0, pool_end = 0;                         // warp-uniform
  while (true) {
    // all lanes converged: rays that are complete wait here; once enough lanes wait they are completed together and idle lanes
    // are refilled together
    const bool fin = busy && L.cur == SENTINEL && L.qn == 0u;
    if (__popc(__ballot_sync(FULL, !busy || fin)) >= REFILL_MIN) {
      if (fin) busy = job.finish(L);
      unsigned idle = __ballot_sync(FULL, !busy);
      while (idle) {
        if (pool_next >= pool_end) {
          if (exhausted) break;
          uint32_t b = 0;
          if (lane == 0) b = atomicAdd(counter, BATCH);
          b = __shfl_sync(FULL, b, 0);
          if (b >= total) { exhausted = true; break; }
          pool_next = b; pool_end = min(b + BATCH, total);
        }
        const uint32_t avail = pool_end - pool_next;
        const uint32_t rank = __popc(idle & lt_mask);
        if (!busy && rank < avail) busy = job.start(pool_next + rank, L);
        pool_next += min(avail, (uint32_t)__popc(idle));
        idle = __ballot_sync(FULL, !busy);
      }
      if (idle == FULL) break;                                  // nothing left anywhere
    }
#pragma unroll 1
    for (int k = 0; k < WINDOW; k++) {
      const bool walking = busy && L.cur != SENTINEL;
      if (walking) lane_visit<STATS, ANYHIT>(L, wnodes, s_top, tris, n, st, S);
      const unsigned queued = __ballot_sync(FULL, L.qn != 0u);
      if (queued) {
        const bool go = __popc(queued) >= FLUSH_MIN || __any_sync(FULL, L.qn >= QFULL) || !__any_sync(FULL, busy && L.cur != SENTINEL);
        if (go) flush_one<STATS, ANYHIT>(L, tris, st);
      }
    }
  }
}

__device__ __forceinline__ void write_contact(clrt_contact* __restrict__ out, uint32_t slot, uint32_t pixelIndex, const Hit& h,
                                              const clrt_triangle* __restrict__ tris, uint32_t pad1 = 0u) {
  uint4 head = make_uint4(pixelIndex, 0u, SENTINEL, pad1);
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

// ---- units -----------------------------------------------------------------------------------------------------------------
// Primary rays: unit u = slot of the (padded) stripe: local tile lt = u / 256, inside the 32x8 tile eight 8x4 sub-blocks of
// 32 consecutive units (so the lanes of a warp start on a compact pixel block); local tile lt of this rank is global
// tile rank + lt*nranks.  raster != 0: contact goes to out[pixelIndex] (the reference's output[c.pixelIndex] = c);
// raster == 0: stripe order, slot = lt*256 + y*32 + x, out-of-image slots get pixelIndex = UINT_MAX.
struct PixelRef { uint32_t pixel, slot; bool inside; };
__device__ __forceinline__ PixelRef pixel_of_unit(const clrt_camera& cam, const TileGrid& g, uint32_t u, uint32_t rank, uint32_t nranks, int raster) {
  const uint32_t lt = u >> 8, sub = (u >> 5) & 7u, l = u & 31u;
#ifdef CLRT_PROBE_TILE_BLOCK      /* experiment only (timing probe): block-cyclic ownership, CLRT_PROBE_TILE_BLOCK consecutive tiles per block */
  const uint32_t gt = (lt / CLRT_PROBE_TILE_BLOCK) * nranks * CLRT_PROBE_TILE_BLOCK + rank * CLRT_PROBE_TILE_BLOCK + lt % CLRT_PROBE_TILE_BLOCK;
#else
  const uint32_t gt = rank + lt * nranks;
#endif
  const uint32_t tx = (gt % g.tilesX) * CLRT_TILE_W + (sub & 3u) * 8u + (l & 7u);
  const uint32_t ty = (gt / g.tilesX) * CLRT_TILE_H + (sub >> 2) * 4u + (l >> 3);
  PixelRef r;
  r.inside = tx < cam.resX && ty < cam.resY && gt < g.tiles;
  r.pixel = ty * cam.resX + tx;
  r.slot = raster ? r.pixel : lt * CLRT_TILE_PIXELS + ((sub >> 2) * 4u + (l >> 3)) * CLRT_TILE_W + (sub & 3u) * 8u + (l & 7u);
  return r;
}

template <bool STATS>
struct PrimaryJob {
  const clrt_camera& cam; TileGrid g; uint32_t rank, nranks; int raster;
  const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; clrt_contact* __restrict__ out; Stats& st;
  uint32_t pixel, slot;
  __device__ __forceinline__ bool start(uint32_t u, Lane& L) {
    const PixelRef r = pixel_of_unit(cam, g, u, rank, nranks, raster);
    if (!r.inside) {
      if (!raster) { Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; write_contact(out, r.slot, SENTINEL, h, tris); }
      return false;
    }
    pixel = r.pixel; slot = r.slot;
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    lane_begin<STATS, 0>(L, tris, n, qf, o, d, RT_FLT_MAX, st);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) { write_contact(out, slot, pixel, L.h, tris, STATS ? L.visits : 0u); return false; }
};

// Generic rays (generateContacts2): contact i for ray i, pixelIndex = i.  A ray with a zero direction cannot hit
// anything in the reference either (det == 0 for every triangle); it is answered without traversal.
template <bool STATS>
struct RaysJob {
  const clrt_ray* __restrict__ rays; const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; clrt_contact* __restrict__ out; Stats& st;
  uint32_t index;
  __device__ __forceinline__ bool start(uint32_t i, Lane& L) {
    const float4* p = reinterpret_cast<const float4*>(rays + i);
    const float4 ro = __ldcs(p + 1), rd = __ldcs(p + 2);
    if (rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f) { Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; write_contact(out, i, i, h, tris); return false; }
    index = i;
    lane_begin<STATS, 0>(L, tris, n, qf, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), RT_FLT_MAX, st);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) { write_contact(out, index, index, L.h, tris); return false; }
};

// Shadow pass + shading fused (ours; the reference shades on the host and has no shadow rays): unit = primary contact;
// for every light an occlusion query (ANYHIT) along the same shadow ray the stand-alone path builds (shade_math.cuh),
// then Blinn-Phong -> RGBA8.  Same pixels as shadow_rays_kernel + trace_rays_kernel + shade_kernel, without the 48-byte
// rays and 32-byte shadow contacts ever touching HBM.  Needs L+1 <= 32 lights.
template <bool STATS>
struct ShadowJob {
  const clrt_camera& cam; const clrt_contact* __restrict__ contacts;
  const clrt_light* __restrict__ lights; uint32_t n_lights; const clrt_material* __restrict__ materials; uint32_t n_materials;
  const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; uint32_t* __restrict__ rgba8; const PeerTargets& peers; Stats& st; uint32_t& live_rays;
  // Per-lane state of a unit: five registers.  The contact's 32 bytes are read again when the unit is shaded and the primary ray
  // is regenerated per shadow ray (40 instructions against ~9000 of a shadow ray's traversal): the 17 registers the whole contact,
  // origin and hit point took cost the kernel a fifth of its occupancy (86 registers, 5 CTAs per SM).
  uint32_t index, pixel, l, occl; float t;
  // peers.n == 0: the local stripe buffer at the contact's slot.  Otherwise the fused exchange: the pixel goes straight into the
  // raster-ordered framebuffer of EVERY GPU of the box (peer memory over NVLink; the local one is among them) — no all-gather,
  // no untile pass.  Slots outside the image (pixel == SENTINEL) have no raster position.
  __device__ __forceinline__ void store_pixel(uint32_t slot, uint32_t px, uint32_t packed) {
    if (peers.n == 0) { __stcs(rgba8 + slot, packed); return; }
    if (px == SENTINEL) return;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) if (k < peers.n) peers.p[k][px] = packed;
  }
  __device__ __forceinline__ void shade_and_store() {
    uint32_t packed;
    const uint32_t oc = occl;
    const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + index));
    const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + index) + 1);
    shade_contact(cam, head, nd, lights, n_lights, materials, n_materials, [oc](uint32_t k, float) { return ((oc >> k) & 1u) != 0u; }, packed);
    store_pixel(index, pixel, packed);
  }
  __device__ __forceinline__ void aim(Lane& L) {
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    const Vec3 P = vadd(o, vscale(d, t));
    float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
    const float limit = vlength(vsub(lp, P)) - SHADOW_EPS_REL * t;
    Vec3 org, dir; shadow_ray(cam, pixel, t, lights, n_lights, l, org, dir);
    lane_begin<STATS, 1>(L, tris, n, qf, org, dir, limit, st);
  }
  __device__ __forceinline__ bool start(uint32_t i, Lane& L) {
    index = i; occl = 0; l = 0;
    const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + i));
    const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + i) + 1);
    pixel = head.x; t = nd.w;
    if (!(nd.w > 0.0f && head.x != SENTINEL)) { store_pixel(i, head.x, 0u); return false; }
    live_rays += n_lights + 1;
    aim(L);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) {
    occl |= (L.h.tri != SENTINEL ? 1u : 0u) << l;
    if (++l <= n_lights) { aim(L); return true; }
    shade_and_store();
    return false;
  }
};

// ---- frame batches ------------------------------------------------------------------------------------------------------------
// Several frames' stripes in ONE launch per pass (multi-GPU: at 8 GPUs a stripe is an eighth of a frame, a launch lives ~0.3 ms and
// what is lost is its drain — the last long rays alone on the machine; four frames per launch make it a half-frame launch).  Unit u =
// frame u / units, stripe slot u % units.  Each frame has its own camera and its own framebuffers (PeerBatch).
constexpr int FRAME_BATCH_MAX = 4;
struct CamBatch { clrt_camera cam[FRAME_BATCH_MAX]; uint32_t frames, units; };        // units: padded stripe slots of one frame
struct PeerBatch { uint32_t* p[FRAME_BATCH_MAX][8]; uint32_t n; };                    // [frame][gpu] raster-ordered framebuffers

template <bool STATS>
struct PrimaryBatchJob {
  const CamBatch& cams; TileGrid g; uint32_t rank, nranks;
  const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; clrt_contact* __restrict__ out; Stats& st;
  uint32_t pixel, slot;
  __device__ __forceinline__ bool start(uint32_t u, Lane& L) {
    const uint32_t f = u / cams.units, u1 = u - f * cams.units;
    const clrt_camera& cam = cams.cam[f];
    const PixelRef r = pixel_of_unit(cam, g, u1, rank, nranks, 0);
    if (!r.inside) { Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; write_contact(out, f * cams.units + r.slot, SENTINEL, h, tris); return false; }
    pixel = r.pixel; slot = f * cams.units + r.slot;
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    lane_begin<STATS, 0>(L, tris, n, qf, o, d, RT_FLT_MAX, st);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) { write_contact(out, slot, pixel, L.h, tris, STATS ? L.visits : 0u); return false; }
};

template <bool STATS>
struct ShadowBatchJob {
  const CamBatch& cams; const clrt_contact* __restrict__ contacts;
  const clrt_light* __restrict__ lights; uint32_t n_lights; const clrt_material* __restrict__ materials; uint32_t n_materials;
  const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; uint32_t* __restrict__ rgba8; const PeerBatch& peers; Stats& st; uint32_t& live_rays;
  uint32_t index, pixel, l, occl; float t;
  __device__ __forceinline__ void store_pixel(uint32_t i, uint32_t px, uint32_t packed) {
    if (peers.n == 0) { __stcs(rgba8 + i, packed); return; }
    if (px == SENTINEL) return;
    const uint32_t f = i / cams.units;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) if (k < peers.n) peers.p[f][k][px] = packed;
  }
  __device__ __forceinline__ void aim(Lane& L) {
    const clrt_camera& cam = cams.cam[index / cams.units];
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    const Vec3 P = vadd(o, vscale(d, t));
    float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
    const float limit = vlength(vsub(lp, P)) - SHADOW_EPS_REL * t;
    Vec3 org, dir; shadow_ray(cam, pixel, t, lights, n_lights, l, org, dir);
    lane_begin<STATS, 1>(L, tris, n, qf, org, dir, limit, st);
  }
  __device__ __forceinline__ bool start(uint32_t i, Lane& L) {
    index = i; occl = 0; l = 0;
    const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + i));
    const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + i) + 1);
    pixel = head.x; t = nd.w;
    if (!(nd.w > 0.0f && head.x != SENTINEL)) { store_pixel(i, head.x, 0u); return false; }
    live_rays += n_lights + 1;
    aim(L);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) {
    occl |= (L.h.tri != SENTINEL ? 1u : 0u) << l;
    if (++l <= n_lights) { aim(L); return true; }
    uint32_t packed;
    const uint32_t oc = occl;
    const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + index));
    const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + index) + 1);
    shade_contact(cams.cam[index / cams.units], head, nd, lights, n_lights, materials, n_materials, [oc](uint32_t k, float) { return ((oc >> k) & 1u) != 0u; }, packed);
    store_pixel(index, pixel, packed);
    return false;
  }
};

// The whole frame in ONE persistent kernel (ours): unit = pixel slot of the stripe; its primary ray (closest hit), then — when it
// hit — one occlusion query per light from the hit point, then Blinn-Phong -> RGBA8.  Same rays, same arithmetic and same pixels as
// trace_primary_kernel + shadow_shade_kernel, without the 32-byte contacts going through memory and without the second kernel's
// launch and drain: a frame (or a 1/8 stripe of one at 8 GPUs) has one tail instead of two.  Lanes of a warp are in either phase.
template <bool STATS>
struct FrameJob {
  const clrt_camera& cam; TileGrid g; uint32_t rank, nranks;
  const clrt_light* __restrict__ lights; uint32_t n_lights; const clrt_material* __restrict__ materials; uint32_t n_materials;
  const clrt_triangle* __restrict__ tris; uint32_t n; const QFrame* __restrict__ qf; uint32_t* __restrict__ rgba8; const PeerTargets& peers;
  Stats& st; uint32_t& live_rays;
  uint32_t pixel, slot, l, occl, tri; float t;      // l == 0: the primary ray is on its way; l >= 1: shadow ray of light l - 1
  __device__ __forceinline__ void store_pixel(uint32_t sl, uint32_t px, uint32_t packed) {
    if (peers.n == 0) { __stcs(rgba8 + sl, packed); return; }
    if (px == SENTINEL) return;
#pragma unroll
    for (uint32_t k = 0; k < 8; k++) if (k < peers.n) peers.p[k][px] = packed;
  }
  __device__ __forceinline__ void aim(Lane& L) {
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    const Vec3 P = vadd(o, vscale(d, t));
    float energy; const Vec3 lp = light_position(lights, n_lights, l - 1u, o, energy);
    const float limit = vlength(vsub(lp, P)) - SHADOW_EPS_REL * t;
    Vec3 org, dir; shadow_ray(cam, pixel, t, lights, n_lights, l - 1u, org, dir);
    L.any = true;
    lane_begin<STATS, 2>(L, tris, n, qf, org, dir, limit, st);
  }
  __device__ __forceinline__ bool start(uint32_t u, Lane& L) {
    const PixelRef r = pixel_of_unit(cam, g, u, rank, nranks, 0);
    if (!r.inside) { store_pixel(r.slot, SENTINEL, 0u); return false; }
    pixel = r.pixel; slot = r.slot; l = 0;
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    L.any = false;
    lane_begin<STATS, 2>(L, tris, n, qf, o, d, RT_FLT_MAX, st);
    return true;
  }
  __device__ __forceinline__ bool finish(Lane& L) {
    if (l == 0) {                                          // the primary ray is complete
      if (L.h.tri == SENTINEL) { store_pixel(slot, pixel, 0u); return false; }
      tri = L.h.tri; t = L.h.t; occl = 0; l = 1;
      live_rays += n_lights + 1;
      aim(L);
      return true;
    }
    occl |= (L.h.tri != SENTINEL ? 1u : 0u) << (l - 1u);
    if (++l <= n_lights + 1u) { aim(L); return true; }
    // shade: the contact write_contact would have stored (normal of the winning triangle, material, t)
    const float4* p = reinterpret_cast<const float4*>(tris + tri);
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    const Vec3 nrm = triangle_normal(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z));
    const uint4 head = make_uint4(pixel, __float_as_uint(a.w), __float_as_uint(b.w), 0u);
    const float4 nd = make_float4(nrm.x, nrm.y, nrm.z, t);
    uint32_t packed;
    const uint32_t oc = occl;
    shade_contact(cam, head, nd, lights, n_lights, materials, n_materials, [oc](uint32_t k, float) { return ((oc >> k) & 1u) != 0u; }, packed);
    store_pixel(slot, pixel, packed);
    return false;
  }
};

// ---- kernels ---------------------------------------------------------------------------------------------------------------
template <int MODE, bool STATS>      // MODE 0: REFERENCE_ORDER, 1: FAST
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
trace_primary_kernel(const clrt_camera cam, const TraceNode* __restrict__ tnodes, const WideNode* __restrict__ wnodes,
                     const clrt_triangle* __restrict__ tris, uint32_t n, clrt_contact* __restrict__ out,
                     uint32_t rank, uint32_t nranks, int raster, uint32_t* __restrict__ counter,
                     unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const uint32_t total = stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  if (MODE == 0) {
    const int lane = threadIdx.x & 31;
    while (true) {                                               // 32-unit packets, one ray per thread, the reference's order
      uint32_t pk = 0;
      if (lane == 0) pk = atomicAdd(counter, 32u);
      pk = __shfl_sync(0xffffffffu, pk, 0);
      if (pk >= total) break;
      const PixelRef r = pixel_of_unit(cam, g, pk + lane, rank, nranks, raster);
      if (r.inside) {
        Vec3 o, d; generate_ray(cam, r.pixel, o, d);
        const Hit h = traverse_ref<STATS>(tnodes, tris, n, o, d, st, overflow);
        write_contact(out, r.slot, r.pixel, h, tris);
      } else if (!raster) {
        Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL;
        write_contact(out, r.slot, SENTINEL, h, tris);
      }
    }
  } else {
    PrimaryJob<STATS> job{cam, g, rank, nranks, raster, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES), out, st, 0u, 0u};
    run_jobs<PrimaryJob<STATS>, STATS, 0>(job, wnodes, tris, n, total, counter, st, overflow);
  }
  if (STATS) flush_stats(st, stats);
}

template <int MODE, bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
trace_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const TraceNode* __restrict__ tnodes, const WideNode* __restrict__ wnodes,
                  const clrt_triangle* __restrict__ tris, uint32_t n, clrt_contact* __restrict__ out,
                  uint32_t* __restrict__ counter, unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  if (MODE == 0) {
    const int lane = threadIdx.x & 31;
    while (true) {
      uint32_t pk = 0;
      if (lane == 0) pk = atomicAdd(counter, 32u);
      pk = __shfl_sync(0xffffffffu, pk, 0);
      if (pk >= count) break;
      const uint32_t i = pk + lane;
      if (i < count) {
        const float4* p = reinterpret_cast<const float4*>(rays + i);
        const float4 ro = __ldcs(p + 1), rd = __ldcs(p + 2);
        Hit h; h.t = RT_FLT_MAX; h.tri = SENTINEL; h.cull = RT_FLT_MAX;
        if (!(rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f))
          h = traverse_ref<STATS>(tnodes, tris, n, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), st, overflow);
        write_contact(out, i, i, h, tris);
      }
    }
  } else {
    RaysJob<STATS> job{rays, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES), out, st, 0u};
    run_jobs<RaysJob<STATS>, STATS, 0>(job, wnodes, tris, n, count, counter, st, overflow);
  }
  if (STATS) flush_stats(st, stats);
}

template <bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
shadow_shade_kernel(const clrt_camera cam, const clrt_contact* __restrict__ contacts, uint32_t count,
                    const clrt_light* __restrict__ lights, uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
                    const WideNode* __restrict__ wnodes, const clrt_triangle* __restrict__ tris, uint32_t n,
                    uint32_t* __restrict__ rgba8, const PeerTargets peers, uint32_t* __restrict__ counter, unsigned long long* __restrict__ ray_counter,
                    unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  uint32_t live_rays = 0;
  ShadowJob<STATS> job{cam, contacts, lights, n_lights, materials, n_materials, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES), rgba8, peers, st, live_rays, 0u, 0u, 0u, 0u, 0.0f};
  run_jobs<ShadowJob<STATS>, STATS, 1>(job, wnodes, tris, n, count, counter, st, overflow);
  if (ray_counter) {
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) live_rays += __shfl_xor_sync(0xffffffffu, live_rays, dd);
    if ((threadIdx.x & 31) == 0 && live_rays) atomicAdd(ray_counter, (unsigned long long)live_rays);
  }
  if (STATS) flush_stats(st, stats);
}

template <bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
frame_kernel(const clrt_camera cam, uint32_t rank, uint32_t nranks,
             const clrt_light* __restrict__ lights, uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
             const WideNode* __restrict__ wnodes, const clrt_triangle* __restrict__ tris, uint32_t n,
             uint32_t* __restrict__ rgba8, const PeerTargets peers, uint32_t* __restrict__ counter, unsigned long long* __restrict__ ray_counter,
             unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  uint32_t live_rays = 0;
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const uint32_t total = stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  FrameJob<STATS> job{cam, g, rank, nranks, lights, n_lights, materials, n_materials, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES),
                      rgba8, peers, st, live_rays, 0u, 0u, 0u, 0u, 0u, 0.0f};
  run_jobs<FrameJob<STATS>, STATS, 2>(job, wnodes, tris, n, total, counter, st, overflow);
  if (ray_counter) {
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) live_rays += __shfl_xor_sync(0xffffffffu, live_rays, dd);
    if ((threadIdx.x & 31) == 0 && live_rays) atomicAdd(ray_counter, (unsigned long long)live_rays);
  }
  if (STATS) flush_stats(st, stats);
}

template <bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
trace_primary_batch_kernel(const CamBatch cams, const WideNode* __restrict__ wnodes, const clrt_triangle* __restrict__ tris, uint32_t n,
                           clrt_contact* __restrict__ out, uint32_t rank, uint32_t nranks, uint32_t* __restrict__ counter,
                           unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  const TileGrid g = make_tile_grid(cams.cam[0].resX, cams.cam[0].resY);
  PrimaryBatchJob<STATS> job{cams, g, rank, nranks, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES), out, st, 0u, 0u};
  run_jobs<PrimaryBatchJob<STATS>, STATS, 0>(job, wnodes, tris, n, cams.frames * cams.units, counter, st, overflow);
  if (STATS) flush_stats(st, stats);
}

template <bool STATS>
__global__ void __launch_bounds__(TRACE_THREADS, CLRT_MIN_CTAS)
shadow_shade_batch_kernel(const CamBatch cams, const clrt_contact* __restrict__ contacts,
                          const clrt_light* __restrict__ lights, uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
                          const WideNode* __restrict__ wnodes, const clrt_triangle* __restrict__ tris, uint32_t n,
                          uint32_t* __restrict__ rgba8, const PeerBatch peers, uint32_t* __restrict__ counter, unsigned long long* __restrict__ ray_counter,
                          unsigned long long* __restrict__ stats, uint32_t* __restrict__ overflow) {
  Stats st; st.rays = st.inner = st.tris = st.lboxes = 0;
  uint32_t live_rays = 0;
  ShadowBatchJob<STATS> job{cams, contacts, lights, n_lights, materials, n_materials, tris, n, reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES),
                            rgba8, peers, st, live_rays, 0u, 0u, 0u, 0u, 0.0f};
  run_jobs<ShadowBatchJob<STATS>, STATS, 1>(job, wnodes, tris, n, cams.frames * cams.units, counter, st, overflow);
  if (ray_counter) {
#pragma unroll
    for (int dd = 16; dd > 0; dd >>= 1) live_rays += __shfl_xor_sync(0xffffffffu, live_rays, dd);
    if ((threadIdx.x & 31) == 0 && live_rays) atomicAdd(ray_counter, (unsigned long long)live_rays);
  }
  if (STATS) flush_stats(st, stats);
}

// Grid of a persistent launch.  Resident capacity (SMs x occupancy, queried once per kernel) is the upper bound.  A launch alone on
// the GPU wants the shortest kernel: one 32-unit packet per warp is enough to start every warp.  A launch in a FRAME SLOT (D
// frames in flight on their own streams) wants throughput, and a warp is efficient only while it lives long: after its last packet
// its lanes drain one by one.  Measured on 1/8 stripes of Soup-1M (profiles/overlap_probe.py): 4 units per lane 81 % of the
// full-frame rate, 16 units per lane with 8 slots 94 %, but 24+ units per lane leave too few warps to fill the machine (86 %, 75 %).
// So: `rays_per_lane` (16) units per lane, but never fewer warps than capacity / D (D kernels run side by side; D = the number of
// frame slots the caller has used so far).
// FAST kernels take FAST_SMEM_BYTES of dynamic shared memory (top table prefix + stack columns); REFERENCE_ORDER kernels none.
template <auto kernel>                         // the kernel itself, not its type: the statics below are per kernel instantiation
int persistent_grid(clrt_accel* a, int threads, size_t smem, uint32_t packets, unsigned* grid) {
  static int per_sm = 0;
  if (per_sm == 0) {
    if (smem > 48 * 1024) CLRT_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, threads, smem));
    if (per_sm < 1) CLRT_FAIL("trace kernel does not fit an SM (%zu bytes of shared memory, %d threads)", smem, threads);
  }
  static int rays_per_lane = 0;
  if (rays_per_lane == 0) { const char* e = getenv("CLRT_RAYS_PER_LANE"); rays_per_lane = e ? atoi(e) : 16; if (rays_per_lane < 1) rays_per_lane = 1; }
  const unsigned resident = (unsigned)(a->ctx->sm_count * per_sm);
  const unsigned warps_per_cta = (unsigned)threads / 32u;
  const unsigned all = (packets + warps_per_cta - 1) / warps_per_cta;               // one packet per warp
  unsigned want = all;
  if (a->launch_stream) {
    const unsigned slots = (unsigned)max(1, a->slots_seen / max(1, a->launch_batch));
    const unsigned fill = (resident + slots - 1u) / slots;
    want = min(all, max((packets + warps_per_cta * rays_per_lane - 1) / (warps_per_cta * rays_per_lane), fill));
  }
  *grid = min(resident, max(1u, want));
  return 0;
}

} // namespace

// The structure a launch traverses: REFERENCE_ORDER always walks the reference topology; FAST walks the separate trace tree
// (trace_tree.cu) when construct() built one, else the reference tree.
// Where a launch goes: the context's stream and the accel's work counter, unless a frame slot lent its own (accel.cu:
// frames in flight on several streams, each with a private counter and scratch).
static inline cudaStream_t trace_stream(const clrt_accel* a) { return a->launch_stream ? a->launch_stream : a->ctx->stream; }
static inline uint32_t* trace_counter(const clrt_accel* a) { return a->launch_counter ? a->launch_counter : a->d_tile_counter; }
static inline uint32_t* overflow_word(const clrt_accel* a) { return a->d_tile_counter + 13; }      // sticky; read by clrt_accel_check

struct TraceView { const TraceNode* tnodes; const WideNode* wnodes; const clrt_triangle* tris; };
static TraceView trace_view(const clrt_accel* a, bool fast) {
  if (fast && a->tt_built) return TraceView{a->d_tt_nodes, a->d_wnodes, a->d_tt_tris};
  return TraceView{a->d_tnodes, a->d_wnodes, a->d_tris_sorted};
}

template <int M, bool S>
static int launch_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks, int raster, uint32_t packets) {
  unsigned grid = 0;
  const size_t smem = M ? FAST_SMEM_BYTES : 0;
  CLRT_TRY((persistent_grid<trace_primary_kernel<M, S>>(a, TRACE_THREADS, smem, packets, &grid)));
  const TraceView v = trace_view(a, M != 0);
  trace_primary_kernel<M, S><<<grid, TRACE_THREADS, smem, trace_stream(a)>>>(cam, v.tnodes, v.wnodes, v.tris, a->n, out, rank, nranks, raster,
                                                                           trace_counter(a), (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  CLRT_CHECK_LAUNCH();
  return 0;
}
template <int M, bool S>
static int launch_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count, uint32_t packets) {
  unsigned grid = 0;
  const size_t smem = M ? FAST_SMEM_BYTES : 0;
  CLRT_TRY((persistent_grid<trace_rays_kernel<M, S>>(a, TRACE_THREADS, smem, packets, &grid)));
  const TraceView v = trace_view(a, M != 0);
  trace_rays_kernel<M, S><<<grid, TRACE_THREADS, smem, trace_stream(a)>>>(rays, count, v.tnodes, v.wnodes, v.tris, a->n, out,
                                                                        trace_counter(a), (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  CLRT_CHECK_LAUNCH();
  return 0;
}
static int trace_mode_index(const clrt_accel* a) { return a->trace_mode == CLRT_TRACE_REFERENCE_ORDER ? 0 : 1; }

int bvh_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks) {
  const uint32_t resX = cam.resX, resY = cam.resY;
  if (!a->built) CLRT_FAIL("generateContacts: the BVH was not constructed");
  const bool raster = (nranks == 0);                 // nranks == 0: whole frame, raster-ordered output
  if (raster) { rank = 0; nranks = 1; }
  const TileGrid g = make_tile_grid(resX, resY);
  const uint32_t packets = stripe_tile_count(g.tiles, rank, nranks) * 8u;
  if (!packets) return 0;
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  const int m = trace_mode_index(a);
  if (a->stats) return m == 0 ? launch_primary<0, true>(a, cam, out, rank, nranks, raster, packets)
                              : launch_primary<1, true>(a, cam, out, rank, nranks, raster, packets);
  return m == 0 ? launch_primary<0, false>(a, cam, out, rank, nranks, raster, packets)
                : launch_primary<1, false>(a, cam, out, rank, nranks, raster, packets);
}

int bvh_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count) {
  if (!a->built) CLRT_FAIL("generateContacts: the BVH was not constructed");
  if (!count) return 0;
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  const int m = trace_mode_index(a);
  if (a->stats) return m == 0 ? launch_rays<0, true>(a, rays, out, count, packets) : launch_rays<1, true>(a, rays, out, count, packets);
  return m == 0 ? launch_rays<0, false>(a, rays, out, count, packets) : launch_rays<1, false>(a, rays, out, count, packets);
}

template <bool S>
static int launch_shadow_shade(clrt_accel* a, const clrt_camera& cam, const clrt_contact* contacts, uint32_t count, uint32_t* rgba8,
                               unsigned long long* d_ray_counter, uint32_t packets) {
  const clrt_scene* sc = a->scene;
  unsigned grid = 0;
  CLRT_TRY(persistent_grid<shadow_shade_kernel<S>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid));
  const TraceView v = trace_view(a, true);
  shadow_shade_kernel<S><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cam, contacts, count, sc->d_lights, sc->n_lights, sc->d_materials, sc->n_materials,
      v.wnodes, v.tris, a->n, rgba8, a->launch_targets, trace_counter(a), d_ray_counter, (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  CLRT_CHECK_LAUNCH();
  return 0;
}

// Fused shadow + shade over `count` contacts (BVH only).  d_ray_counter (may be NULL) accumulates the live shadow rays.
int bvh_shadow_shade(clrt_accel* a, const clrt_camera& cam, const clrt_contact* contacts, uint32_t count, uint32_t* rgba8,
                     unsigned long long* d_ray_counter) {
  if (!a->built) CLRT_FAIL("shadow pass: the BVH was not constructed");
  if (!count) return 0;
  if (a->scene->n_lights + 1 > 32) CLRT_FAIL("fused shadow pass supports at most 31 scene lights");
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  return a->stats ? launch_shadow_shade<true>(a, cam, contacts, count, rgba8, d_ray_counter, packets)
                  : launch_shadow_shade<false>(a, cam, contacts, count, rgba8, d_ray_counter, packets);
}

template <bool S>
static int launch_frame(clrt_accel* a, const clrt_camera& cam, uint32_t rank, uint32_t nranks, uint32_t* rgba8, unsigned long long* d_ray_counter, uint32_t packets) {
  const clrt_scene* sc = a->scene;
  unsigned grid = 0;
  CLRT_TRY(persistent_grid<frame_kernel<S>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid));
  const TraceView v = trace_view(a, true);
  frame_kernel<S><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cam, rank, nranks, sc->d_lights, sc->n_lights, sc->d_materials, sc->n_materials,
      v.wnodes, v.tris, a->n, rgba8, a->launch_targets, trace_counter(a), d_ray_counter, (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  CLRT_CHECK_LAUNCH();
  return 0;
}

// One stripe of a frame — primary rays, shadow rays, shading — in one launch (BVH, FAST policy; stripe-ordered RGBA8 or, with peer
// targets, raster-ordered stores into every framebuffer).  d_ray_counter (may be NULL) accumulates the live shadow rays.
int bvh_frame_fused(clrt_accel* a, const clrt_camera& cam, uint32_t rank, uint32_t nranks, uint32_t* rgba8, unsigned long long* d_ray_counter) {
  if (!a->built) CLRT_FAIL("frame: the BVH was not constructed");
  if (a->scene->n_lights + 1 > 32) CLRT_FAIL("fused frame supports at most 31 scene lights");
  const TileGrid g = make_tile_grid(cam.resX, cam.resY);
  const uint32_t packets = stripe_tile_count(g.tiles, rank, nranks) * 8u;
  if (!packets) return 0;
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  return a->stats ? launch_frame<true>(a, cam, rank, nranks, rgba8, d_ray_counter, packets)
                  : launch_frame<false>(a, cam, rank, nranks, rgba8, d_ray_counter, packets);
}

// `frames` (1..4) stripes `rank` of `nranks` — one camera each, same resolution — as two launches: primary contacts of all of them
// into `contacts` (frames x units), then shadow + shade into `rgba8` (frames x units, stripe order) or, with peer targets
// (targets[f * 8 + gpu], n_targets GPUs), into the raster framebuffers of frame f on every GPU.  FAST policy.
int bvh_frames_batch(clrt_accel* a, const clrt_camera* cams, uint32_t frames, uint32_t rank, uint32_t nranks, clrt_contact* contacts,
                     uint32_t* rgba8, uint32_t* const* targets, uint32_t n_targets, unsigned long long* d_ray_counter) {
  if (!a->built) CLRT_FAIL("frame batch: the BVH was not constructed");
  if (!frames || frames > (uint32_t)FRAME_BATCH_MAX) CLRT_FAIL("frame batch: %u frames (1..%d)", frames, FRAME_BATCH_MAX);
  if (a->scene->n_lights + 1 > 32) CLRT_FAIL("fused shadow pass supports at most 31 scene lights");
  CamBatch cb; memset(&cb, 0, sizeof cb);
  for (uint32_t f = 0; f < frames; f++) {
    cb.cam[f] = cams[f];
    if (cams[f].resX != cams[0].resX || cams[f].resY != cams[0].resY) CLRT_FAIL("frame batch: the cameras differ in resolution");
  }
  const TileGrid g = make_tile_grid(cams[0].resX, cams[0].resY);
  cb.frames = frames; cb.units = stripe_tile_count(g.tiles, rank, nranks) * CLRT_TILE_PIXELS;
  if (!cb.units) return 0;
  PeerBatch pb; memset(&pb, 0, sizeof pb);
  pb.n = n_targets;
  for (uint32_t f = 0; f < frames && n_targets; f++) for (uint32_t k = 0; k < n_targets; k++) pb.p[f][k] = targets[f * 8 + k];
  const uint32_t packets = frames * (cb.units / 32u);
  const clrt_scene* sc = a->scene;
  const TraceView v = trace_view(a, true);
  unsigned grid = 0;
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  if (a->stats) {
    CLRT_TRY((persistent_grid<trace_primary_batch_kernel<true>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid)));
    trace_primary_batch_kernel<true><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cb, v.wnodes, v.tris, a->n, contacts, rank, nranks, trace_counter(a),
        (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  } else {
    CLRT_TRY((persistent_grid<trace_primary_batch_kernel<false>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid)));
    trace_primary_batch_kernel<false><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cb, v.wnodes, v.tris, a->n, contacts, rank, nranks, trace_counter(a),
        (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  }
  CLRT_CHECK_LAUNCH();
  CLRT_CUDA(cudaMemsetAsync(trace_counter(a), 0, 4, trace_stream(a)));
  if (a->stats) {
    CLRT_TRY((persistent_grid<shadow_shade_batch_kernel<true>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid)));
    shadow_shade_batch_kernel<true><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cb, contacts, sc->d_lights, sc->n_lights, sc->d_materials, sc->n_materials,
        v.wnodes, v.tris, a->n, rgba8, pb, trace_counter(a), d_ray_counter, (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  } else {
    CLRT_TRY((persistent_grid<shadow_shade_batch_kernel<false>>(a, TRACE_THREADS, FAST_SMEM_BYTES, packets, &grid)));
    shadow_shade_batch_kernel<false><<<grid, TRACE_THREADS, FAST_SMEM_BYTES, trace_stream(a)>>>(cb, contacts, sc->d_lights, sc->n_lights, sc->d_materials, sc->n_materials,
        v.wnodes, v.tris, a->n, rgba8, pb, trace_counter(a), d_ray_counter, (unsigned long long*)(a->d_tile_counter + 4), overflow_word(a));
  }
  CLRT_CHECK_LAUNCH();
  return 0;
}

// Traversal stacks are bounded (REF_STACK, FAST_STACK).  A push that does not fit is counted on the device instead of silently
// dropping a subtree; this call synchronises the context's stream and FAILS if any launch since the last check overflowed.
extern "C" int clrt_accel_check(clrt_accel* a) {
  if (!a) CLRT_FAIL("clrt_accel_check: NULL accel");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  void* pin = nullptr;
  CLRT_TRY(clrt_ctx_pinned(a->ctx, 64, &pin));
  CLRT_CUDA(cudaMemcpyAsync(pin, a->d_tile_counter + 13, 4, cudaMemcpyDeviceToHost, a->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(a->ctx->stream));
  const uint32_t dropped = *(uint32_t*)pin;
  if (dropped) {
    CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter + 13, 0, 4, a->ctx->stream));
    CLRT_FAIL("traversal stack overflow: %u pushes did not fit (tree deeper than the %d / %d entry stacks); contacts are incomplete", dropped, REF_STACK, FAST_STACK);
  }
  return 0;
}
