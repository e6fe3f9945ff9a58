// kd.cu — KD-tree manager: a third AccelerationStructureManager.
//
// The reference has no KD-tree; its KNOWN_ISSUES_AND_TODO.txt:24-26 asks for one: "Adding KD trees - Construction may be based on
// method which is used to construct the BVH ... Since KD tree traversal is relatively fast, using fast construction algorithm could
// yield better performance".  The plug-in point is AccelerationStructureManager (R/Include/Algorithms/AccelerationStructureManager.h:
// 52-78); this file is the device side behind the same C ABI handle (clrt_accel) and the same clrt_accel_* / clrt_trace_* calls.
//
// Construction = the LBVH's method (Morton codes -> sort -> Karras radix tree, BVH.h:83-321) applied to SPACE instead of objects:
//   1. the scene box is cut into 2^L cells per axis (L chosen so that there are about two cells per triangle, L <= 10);
//   2. every triangle emits one (Morton code of the cell, triangle) pair per cell its box overlaps: count -> scan -> scatter
//      (the Two-Level Grid's pattern, TwoLevelGridManager.cpp:207-345, with the shared scan block);
//   3. the pairs are sorted by code with the shared onesweep sort; runs of equal codes are the LEAVES (triangle lists of occupied cells);
//   4. a Karras radix tree over the UNIQUE codes is a kd-tree: the first bit in which the two halves of a node's range differ names
//      an axis (bits cycle x, y, z from the top) and a level, i.e. an axis-aligned plane through the middle of the node's region.
//      Empty cells cost nothing: they belong to the region of a neighbouring leaf, whose triangle list is complete for them (empty).
// Traversal: the classic kd-tree walk with a [tmin, tmax] interval and a stack of far children; leaves are visited in ray order, the
// closest hit so far is kept, and the walk ends as soon as it lies inside the region just finished.  Result: the true closest hit —
// what the BVH returns (checked against the oracle's BVH contacts by the north-star tolerances, tests/test_gpu_kd.py).
#include "common.cuh"
#include "rtmath.cuh"

struct KdNode { float plane; uint32_t axis; uint32_t childA, childB; };   // child: < m leaf (index into the unique-code array), else m + inner index

struct KdState {
  int levels = 0;                           // L: 2^L cells per axis
  float lo[3] = {0, 0, 0}, cs[3] = {0, 0, 0};   // grid origin and cell size per axis
  uint32_t pairs_cap = 0, pairs = 0, leaves = 0;
  uint32_t* d_counters = nullptr; size_t counters_cap = 0;   // per triangle, later per pair (run heads)
  uint32_t* d_scan = nullptr; size_t scan_cap = 0;
  clrt_uint2* d_pairs = nullptr; size_t pairs_capacity = 0;  // sorted (code, triangle)
  uint32_t* d_codes = nullptr; size_t codes_cap = 0;         // unique codes, ascending
  clrt_uint2* d_ranges = nullptr; size_t ranges_cap = 0;     // per leaf: [first, end) into d_pairs
  KdNode* d_nodes = nullptr; size_t nodes_cap = 0;           // leaves - 1 inner nodes, root = 0
  bool built = false;
};

namespace {

struct KdGrid { float lox, loy, loz, csx, csy, csz; int levels; clrt_aabb box; };   // box: the scene box (tie-break by the reference's order)

__device__ __forceinline__ uint32_t kd_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t kd_code(uint32_t x, uint32_t y, uint32_t z) { return kd_expand_bits(x) * 4 + kd_expand_bits(y) * 2 + kd_expand_bits(z); }

struct CellBox { uint32_t x0, x1, y0, y1, z0, z1; };
__device__ __forceinline__ CellBox kd_cell_box(const clrt_triangle* __restrict__ tris, uint32_t i, const KdGrid& g) {
  const float4* t = reinterpret_cast<const float4*>(tris + i);
  const float4 a = __ldg(t), b = __ldg(t + 1), c = __ldg(t + 2);
  const float mnx = fminf(a.x, fminf(b.x, c.x)), mny = fminf(a.y, fminf(b.y, c.y)), mnz = fminf(a.z, fminf(b.z, c.z));
  const float mxx = fmaxf(a.x, fmaxf(b.x, c.x)), mxy = fmaxf(a.y, fmaxf(b.y, c.y)), mxz = fmaxf(a.z, fmaxf(b.z, c.z));
  const int r = (1 << g.levels) - 1;
  // The walk decides sides by planes (lo + cut * cs), the lists by floor((v - lo) / cs): the two can disagree by rounding (up to
  // ~1e-4 of a cell at 1024 cells per axis).  The box is widened by 1e-3 of a cell, so a triangle that touches a plane within that
  // margin is listed on both sides.
  const float ex = 1.0e-3f * g.csx, ey = 1.0e-3f * g.csy, ez = 1.0e-3f * g.csz;
  auto cell = [r](float v, float lo, float cs) { const int q = (int)floorf((v - lo) / cs); return (uint32_t)(q < 0 ? 0 : (q > r ? r : q)); };
  CellBox cb;
  cb.x0 = cell(mnx - ex, g.lox, g.csx); cb.x1 = cell(mxx + ex, g.lox, g.csx);
  cb.y0 = cell(mny - ey, g.loy, g.csy); cb.y1 = cell(mxy + ey, g.loy, g.csy);
  cb.z0 = cell(mnz - ez, g.loz, g.csz); cb.z1 = cell(mxz + ez, g.loz, g.csz);
  return cb;
}

__global__ void __launch_bounds__(256)
kd_count_kernel(const clrt_triangle* __restrict__ tris, uint32_t n, const KdGrid g, uint32_t* __restrict__ counters) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const CellBox cb = kd_cell_box(tris, i, g);
  counters[i] = (cb.x1 - cb.x0 + 1) * (cb.y1 - cb.y0 + 1) * (cb.z1 - cb.z0 + 1);
}

__global__ void __launch_bounds__(256)
kd_scatter_kernel(const clrt_triangle* __restrict__ tris, uint32_t n, const KdGrid g, const uint32_t* __restrict__ scan,
                  const uint32_t* __restrict__ counters, uint32_t cap, uint2* __restrict__ pairs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const CellBox cb = kd_cell_box(tris, i, g);
  uint32_t at = scan[i] - counters[i];
  for (uint32_t z = cb.z0; z <= cb.z1; z++)
    for (uint32_t y = cb.y0; y <= cb.y1; y++)
      for (uint32_t x = cb.x0; x <= cb.x1; x++, at++)
        if (at < cap) pairs[at] = make_uint2(kd_code(x, y, z), i);
}

// run heads of the sorted pairs -> 1 / 0 (scanned into leaf numbers)
__global__ void __launch_bounds__(256)
kd_heads_kernel(const uint2* __restrict__ pairs, uint32_t count, uint32_t* __restrict__ heads) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  heads[k] = (k == 0 || pairs[k].x != pairs[k - 1].x) ? 1u : 0u;
}
// leaf l = scan[k] - 1 for a run head k: its code, the start of its range; the end is written by the next head (or the last pair)
__global__ void __launch_bounds__(256)
kd_leaves_kernel(const uint2* __restrict__ pairs, uint32_t count, const uint32_t* __restrict__ scan, uint32_t* __restrict__ codes,
                 uint2* __restrict__ ranges) {
  const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= count) return;
  const uint32_t code = pairs[k].x;
  const bool head = k == 0 || code != pairs[k - 1].x, tail = k + 1 == count || code != pairs[k + 1].x;
  const uint32_t leaf = scan[k] - 1u;
  if (head) { codes[leaf] = code; ranges[leaf].x = k; }
  if (tail) ranges[leaf].y = k + 1u;
}

// Karras radix tree over m UNIQUE ascending codes (BVH.h:138-321 without the duplicate-key chains): node i's range, split, and the
// plane the split stands for.  delta(i, j) = common prefix length of codes i and j (-1 outside the array).
__device__ __forceinline__ int kd_delta(const uint32_t* __restrict__ codes, int m, uint32_t ci, int j) {
  if (j < 0 || j >= m) return -1;
  return __clz(ci ^ __ldg(codes + j));
}
__global__ void __launch_bounds__(256)
kd_emit_kernel(const uint32_t* __restrict__ codes, uint32_t m_, const KdGrid g, KdNode* __restrict__ nodes) {
  const int m = (int)m_;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= m) return;
  const uint32_t ci = __ldg(codes + i);
  const int d = kd_delta(codes, m, ci, i + 1) - kd_delta(codes, m, ci, i - 1) >= 0 ? 1 : -1;
  const int dmin = kd_delta(codes, m, ci, i - d);
  int lmax = 2;
  while (kd_delta(codes, m, ci, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (kd_delta(codes, m, ci, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = kd_delta(codes, m, ci, j);
  int s = 0;
  for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
    if (kd_delta(codes, m, ci, i + (s + t) * d) > dnode) s += t;
    if (t <= 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int first = min(i, j), last = max(i, j);
  // The first differing bit: dnode leading bits are common (codes are 32-bit words holding 3L significant bits).  Bit position from the
  // top of the 3L-bit code: p = dnode - (32 - 3L); axis = p % 3 (x, y, z), level = p / 3 (0 = the coarsest cut of that axis).
  const int p = dnode - (32 - 3 * g.levels);
  const int axis = p % 3, level = p / 3;
  // coordinate bits of that axis above the level are common to the whole range: take them from the first code
  const uint32_t cf = __ldg(codes + first);
  uint32_t coord = 0;                                     // compact the axis' bits of cf
  for (int b = 0; b < g.levels; b++) coord |= ((cf >> (3 * b + (2 - axis))) & 1u) << b;
  const int shift = g.levels - 1 - level;                 // the split bit inside the coordinate
  const uint32_t cut = ((coord >> (shift + 1)) << (shift + 1)) | (1u << shift);     // first cell index on the upper side
  const float lo = axis == 0 ? g.lox : axis == 1 ? g.loy : g.loz, cs = axis == 0 ? g.csx : axis == 1 ? g.csy : g.csz;
  KdNode nd;
  nd.plane = lo + (float)cut * cs;
  nd.axis = (uint32_t)axis;
  nd.childA = first == gamma ? (uint32_t)gamma : (uint32_t)gamma + m_;
  nd.childB = last == gamma + 1 ? (uint32_t)gamma + 1u : (uint32_t)gamma + 1u + m_;
  nodes[i] = nd;
}

// ---- traversal ---------------------------------------------------------------------------------------------------------------
struct KdHit { float t; uint32_t tri; };
constexpr int KD_STACK = 40;          // far children: at most one per level, 3L <= 30 levels
constexpr int KD_THREADS = 128;

#ifndef CLRT_KD_REFILL
#define CLRT_KD_REFILL 1             /* 0: kd_trace_*_kernel (one ray per thread to the end) */
#endif
#ifndef CLRT_KD_REFILL_MIN
#define CLRT_KD_REFILL_MIN 12
#endif
#ifndef CLRT_KD_MIN_CTAS
#define CLRT_KD_MIN_CTAS 8
#endif
#if !CLRT_KD_REFILL
__device__ __forceinline__ KdHit kd_traverse(Vec3 o, Vec3 d, const KdGrid& g, const KdNode* __restrict__ nodes, uint32_t m,
                                             const uint2* __restrict__ ranges, const uint2* __restrict__ pairs,
                                             const clrt_triangle* __restrict__ tris, uint32_t* __restrict__ overflow) {
  KdHit best; best.t = RT_FLT_MAX; best.tri = 0xFFFFFFFFu;
  KdHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
  if (m == 0) return none;
  const float big = 1.0e30f;
  const float inv[3] = {d.x == 0.0f ? copysignf(big, d.x) : 1.0f / d.x, d.y == 0.0f ? copysignf(big, d.y) : 1.0f / d.y, d.z == 0.0f ? copysignf(big, d.z) : 1.0f / d.z};
  const float org[3] = {o.x, o.y, o.z}, dir[3] = {d.x, d.y, d.z};
  float tmin, tmax;
  {
    const float ext = (float)(1 << g.levels);
    const float ax = (g.lox - o.x) * inv[0], bx = (g.lox + ext * g.csx - o.x) * inv[0];
    const float ay = (g.loy - o.y) * inv[1], by = (g.loy + ext * g.csy - o.y) * inv[1];
    const float az = (g.loz - o.z) * inv[2], bz = (g.loz + ext * g.csz - o.z) * inv[2];
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz)), tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    if (!(tn <= tf) || tf < 0.0f) return none;
    tmin = fmaxf(tn, 0.0f); tmax = tf;
  }
  // One unit of work per loop iteration and lane — descend one inner node, open a leaf, or test one triangle — so that the lanes of
  // a warp stay in step (nested per-lane loops ran with 4 of 32 lanes: every lane waited for the longest inner loop of the warp).
  uint2 stack[KD_STACK];                                           // (far child, its tmax)
  int sp = 0;
  uint32_t cur = m >= 2 ? m : 0u;                                  // root = inner node 0, or the only leaf
  uint32_t k = 0, kend = 0;
  constexpr uint32_t IN_LEAF = 0xFFFFFFFEu;
  const float slack = 1.00001f;
  while (true) {
    if (cur == IN_LEAF) {
      if (k < kend) {
        const uint32_t id = __ldg(&pairs[k].y); k++;
        const float4* t4 = reinterpret_cast<const float4*>(tris + id);
        const float4 a = __ldg(t4), b = __ldg(t4 + 1), c = __ldg(t4 + 2);
        const float t = triangle_intersect(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), o, d);
        // bit-equal t: the reference's traversal keeps the triangle that comes first in its sorted (Morton code, index) order
        if (t > 0.0f && (t < best.t || (t == best.t && reference_order_before(tris, id, best.tri, g.box)))) { best.t = t; best.tri = id; }
        continue;
      }
      if (best.t <= tmax * slack) return best;                     // the closest hit lies in a region already finished
      if (sp == 0) break;
      sp--;
      tmin = tmax; tmax = __uint_as_float(stack[sp].y); cur = stack[sp].x;
      continue;
    }
    if (cur >= m) {                                                // inner node: pick the near side, remember the far one
      const KdNode nd = nodes[cur - m];
      const float oa = nd.axis == 0 ? org[0] : nd.axis == 1 ? org[1] : org[2];
      const float da = nd.axis == 0 ? dir[0] : nd.axis == 1 ? dir[1] : dir[2];
      const float ia = nd.axis == 0 ? inv[0] : nd.axis == 1 ? inv[1] : inv[2];
      const float ts = (nd.plane - oa) * ia;
      const bool below = oa < nd.plane || (oa == nd.plane && da <= 0.0f);
      const uint32_t near = below ? nd.childA : nd.childB, far = below ? nd.childB : nd.childA;
      if (da == 0.0f || ts > tmax * slack || ts <= 0.0f) cur = near;          // the ray never reaches the plane inside the interval
      else if (ts < tmin / slack) cur = far;                                  // it crossed the plane before the interval
      else {
        if (sp < KD_STACK) { stack[sp] = make_uint2(far, __float_as_uint(tmax)); sp++; } else atomicAdd(overflow, 1u);
        cur = near; tmax = ts;
      }
    } else {                                                       // leaf: open its triangle list
      const uint2 range = __ldg(ranges + cur);
      k = range.x; kend = range.y; cur = IN_LEAF;
    }
  }
  return best.tri != 0xFFFFFFFFu ? best : none;
}

#endif

__device__ __forceinline__ void kd_write_contact(clrt_contact* __restrict__ out, uint32_t slot, uint32_t pixelIndex, const KdHit& h,
                                                 const clrt_triangle* __restrict__ tris) {
  uint4 head = make_uint4(pixelIndex, 0u, 0xFFFFFFFFu, 0u);
  float4 nd = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
  if (h.tri != 0xFFFFFFFFu) {
    const float4* p = reinterpret_cast<const float4*>(tris + h.tri);
    const float4 a = __ldg(p), b = __ldg(p + 1), c = __ldg(p + 2);
    const Vec3 nrm = triangle_normal(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z));
    nd = make_float4(nrm.x, nrm.y, nrm.z, h.t);
    head.y = __float_as_uint(a.w);
    head.z = h.tri;
  }
  uint4* o = reinterpret_cast<uint4*>(out + slot);
  __stcs(o, head);
  __stcs(reinterpret_cast<float4*>(o) + 1, nd);
}

#if !CLRT_KD_REFILL
// generateContacts(Camera&): persistent warps over 8x4-pixel packets (the tiling of the other two structures: stripes of 32x8 tiles)
__global__ void __launch_bounds__(KD_THREADS)
kd_trace_primary_kernel(const clrt_camera cam, const KdGrid g, const KdNode* __restrict__ nodes, uint32_t m, const uint2* __restrict__ ranges,
                        const uint2* __restrict__ pairs, const clrt_triangle* __restrict__ tris, clrt_contact* __restrict__ out,
                        uint32_t rank, uint32_t nranks, int raster, uint32_t* __restrict__ counter, uint32_t* __restrict__ overflow) {
  const TileGrid tg = make_tile_grid(cam.resX, cam.resY);
  const uint32_t packets = stripe_tile_count(tg.tiles, rank, nranks) * 8u;
  const int lane = threadIdx.x & 31;
  while (true) {
    uint32_t pk = 0;
    if (lane == 0) pk = atomicAdd(counter, 1u);
    pk = __shfl_sync(0xffffffffu, pk, 0);
    if (pk >= packets) break;
    const uint32_t lt = pk >> 3, sub = pk & 7u;
    const uint32_t gt = rank + lt * nranks;
    const uint32_t tx = (gt % tg.tilesX) * CLRT_TILE_W + (sub & 3u) * 8u + (lane & 7);
    const uint32_t ty = (gt / tg.tilesX) * CLRT_TILE_H + (sub >> 2) * 4u + (lane >> 3);
    const uint32_t inTile = ((sub >> 2) * 4u + (lane >> 3)) * CLRT_TILE_W + (sub & 3u) * 8u + (lane & 7);
    const bool inside = tx < cam.resX && ty < cam.resY;
    const uint32_t pixel = ty * cam.resX + tx;
    if (inside) {
      Vec3 o, d; generate_ray(cam, pixel, o, d);
      const KdHit h = kd_traverse(o, d, g, nodes, m, ranges, pairs, tris, overflow);
      kd_write_contact(out, raster ? pixel : lt * CLRT_TILE_PIXELS + inTile, pixel, h, tris);
    } else if (!raster) {
      KdHit h; h.t = 0.0f; h.tri = 0xFFFFFFFFu;
      kd_write_contact(out, lt * CLRT_TILE_PIXELS + inTile, 0xFFFFFFFFu, h, tris);
    }
  }
}

// generateContacts(rays, contacts, rayCount)
__global__ void __launch_bounds__(KD_THREADS)
kd_trace_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const KdGrid g, const KdNode* __restrict__ nodes, uint32_t m,
                     const uint2* __restrict__ ranges, const uint2* __restrict__ pairs, const clrt_triangle* __restrict__ tris,
                     clrt_contact* __restrict__ out, uint32_t* __restrict__ counter, uint32_t* __restrict__ overflow) {
  const int lane = threadIdx.x & 31;
  const uint32_t packets = (count + 31u) / 32u;
  while (true) {
    uint32_t pk = 0;
    if (lane == 0) pk = atomicAdd(counter, 1u);
    pk = __shfl_sync(0xffffffffu, pk, 0);
    if (pk >= packets) break;
    const uint32_t i = pk * 32u + lane;
    if (i < count) {
      const float4* p = reinterpret_cast<const float4*>(rays + i);
      const float4 ro = __ldcs(p + 1), rd = __ldcs(p + 2);
      KdHit h; h.t = 0.0f; h.tri = 0xFFFFFFFFu;
      if (!(rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f))
        h = kd_traverse(v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), g, nodes, m, ranges, pairs, tris, overflow);
      kd_write_contact(out, i, i, h, tris);
    }
  }
}

#endif

// ---- the same walk on persistent warps with lane refill ------------------------------------------------------------------------------
// kd_traverse runs one ray per thread as a single-loop state machine (4 of 32 lanes per instruction in ncu: three branch bodies per
// iteration and rays of very different lengths).  What worked for the grid (profiles/r02_experiments.md section 7) is applied here:
// compact nested loops, resumable between LEAVES — one step of a lane = descend to the next leaf, test its triangles, decide, pop —
// and lanes that are complete wait until enough lanes wait, write their contacts and take new rays together.  Same arithmetic and
// order of operations per ray as kd_traverse, hence the same hits.  The stack of far children (local memory) is part of the lane.
struct KdView { KdGrid g; const KdNode* nodes; uint32_t m; const uint2* ranges; const uint2* pairs; const clrt_triangle* tris; uint32_t* overflow; };
struct KLane {
  Vec3 o, d, inv;
  KdHit best;
  float tmin, tmax;
  uint32_t cur;
  int sp;
  bool done;
  uint32_t slot, pixel;
};                                    // the stack of far children is a separate local array: inside the struct it dragged every field into local memory
__device__ __forceinline__ bool klane_begin(KLane& L, const KdView& V, Vec3 o, Vec3 d) {
  const KdGrid& g = V.g;
  L.o = o; L.d = d; L.best.t = RT_FLT_MAX; L.best.tri = 0xFFFFFFFFu; L.done = true;
  if (V.m == 0) return false;
  const float big = 1.0e30f;
  L.inv = v3(d.x == 0.0f ? copysignf(big, d.x) : 1.0f / d.x, d.y == 0.0f ? copysignf(big, d.y) : 1.0f / d.y, d.z == 0.0f ? copysignf(big, d.z) : 1.0f / d.z);
  const float ext = (float)(1 << g.levels);
  const float ax = (g.lox - o.x) * L.inv.x, bx = (g.lox + ext * g.csx - o.x) * L.inv.x;
  const float ay = (g.loy - o.y) * L.inv.y, by = (g.loy + ext * g.csy - o.y) * L.inv.y;
  const float az = (g.loz - o.z) * L.inv.z, bz = (g.loz + ext * g.csz - o.z) * L.inv.z;
  const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz)), tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  if (!(tn <= tf) || tf < 0.0f) return false;
  L.tmin = fmaxf(tn, 0.0f); L.tmax = tf;
  L.sp = 0; L.cur = V.m >= 2 ? V.m : 0u;
  L.done = false;
  return true;
}
__device__ __forceinline__ void klane_leaf_step(KLane& L, uint2* __restrict__ stack, const KdView& V) {
  const float slack = 1.00001f;
  const uint32_t m = V.m;
  uint32_t cur = L.cur;
  float tmin = L.tmin, tmax = L.tmax;
  while (cur >= m) {                                             // inner nodes: pick the near side, remember the far one
    const KdNode nd = V.nodes[cur - m];
    const float oa = nd.axis == 0 ? L.o.x : nd.axis == 1 ? L.o.y : L.o.z;
    const float da = nd.axis == 0 ? L.d.x : nd.axis == 1 ? L.d.y : L.d.z;
    const float ia = nd.axis == 0 ? L.inv.x : nd.axis == 1 ? L.inv.y : L.inv.z;
    const float ts = (nd.plane - oa) * ia;
    const bool below = oa < nd.plane || (oa == nd.plane && da <= 0.0f);
    const uint32_t near = below ? nd.childA : nd.childB, far = below ? nd.childB : nd.childA;
    if (da == 0.0f || ts > tmax * slack || ts <= 0.0f) cur = near;
    else if (ts < tmin / slack) cur = far;
    else {
      if (L.sp < KD_STACK) { stack[L.sp] = make_uint2(far, __float_as_uint(tmax)); L.sp++; } else atomicAdd(V.overflow, 1u);
      cur = near; tmax = ts;
    }
  }
  const uint2 range = __ldg(V.ranges + cur);
  for (uint32_t k = range.x; k < range.y; k++) {
    const uint32_t id = __ldg(&V.pairs[k].y);
    const float4* t4 = reinterpret_cast<const float4*>(V.tris + id);
    const float4 a = __ldg(t4), b = __ldg(t4 + 1), c = __ldg(t4 + 2);
    const float t = triangle_intersect(v3(a.x, a.y, a.z), v3(b.x, b.y, b.z), v3(c.x, c.y, c.z), L.o, L.d);
    if (t > 0.0f && (t < L.best.t || (t == L.best.t && reference_order_before(V.tris, id, L.best.tri, V.g.box)))) { L.best.t = t; L.best.tri = id; }
  }
  if (L.best.t <= tmax * slack || L.sp == 0) { L.done = true; return; }      // the closest hit lies in a region already finished / nothing left
  L.sp--;
  const uint2 e = stack[L.sp];
  L.tmin = tmax; L.tmax = __uint_as_float(e.y); L.cur = e.x;
}

template <class Job>
__device__ __forceinline__ void kd_run_jobs(Job& job, const KdView& V, uint32_t total, uint32_t* __restrict__ counter, clrt_contact* __restrict__ out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  KLane L; L.done = true;
  uint2 stack[KD_STACK];                                           // (far child, its tmax) of the lane's ray
  bool busy = false, exhausted = false;
  uint32_t pool_next = 0, pool_end = 0;
  while (true) {
    const bool fin = busy && L.done;
    const unsigned waiting = __ballot_sync(FULL, !busy || fin);
    if (__popc(waiting) >= CLRT_KD_REFILL_MIN || !__any_sync(FULL, busy && !fin)) {
      if (fin) {
        KdHit h = L.best; if (h.tri == 0xFFFFFFFFu) h.t = 0.0f;
        kd_write_contact(out, L.slot, L.pixel, h, V.tris);
        busy = false;
      }
      unsigned idle = __ballot_sync(FULL, !busy);
      while (idle) {
        if (pool_next >= pool_end) {
          if (exhausted) break;
          uint32_t b = 0;
          if (lane == 0) b = atomicAdd(counter, 32u);
          b = __shfl_sync(FULL, b, 0);
          if (b >= total) { exhausted = true; break; }
          pool_next = b; pool_end = min(b + 32u, total);
        }
        const uint32_t avail = pool_end - pool_next;
        const uint32_t rnk = __popc(idle & lt_mask);
        if (!busy && rnk < avail) busy = job.start(pool_next + rnk, L, V, out);
        pool_next += min(avail, (uint32_t)__popc(idle));
        idle = __ballot_sync(FULL, !busy);
      }
      if (idle == FULL) break;
    }
    if (busy && !L.done) klane_leaf_step(L, stack, V);
  }
}

struct KdPrimaryJob {
  clrt_camera cam; TileGrid tg; uint32_t rank, nranks; int raster;
  __device__ __forceinline__ bool start(uint32_t unit, KLane& L, const KdView& V, clrt_contact* __restrict__ out) const {
    const uint32_t pk = unit >> 5, ln = unit & 31u;
    const uint32_t lt = pk >> 3, sub = pk & 7u;
    const uint32_t gt = rank + lt * nranks;
    const uint32_t tx = (gt % tg.tilesX) * CLRT_TILE_W + (sub & 3u) * 8u + (ln & 7u);
    const uint32_t ty = (gt / tg.tilesX) * CLRT_TILE_H + (sub >> 2) * 4u + (ln >> 3);
    const uint32_t inTile = ((sub >> 2) * 4u + (ln >> 3)) * CLRT_TILE_W + (sub & 3u) * 8u + (ln & 7u);
    const bool inside = tx < cam.resX && ty < cam.resY;
    const uint32_t pixel = ty * cam.resX + tx;
    KdHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
    if (!inside) { if (!raster) kd_write_contact(out, lt * CLRT_TILE_PIXELS + inTile, 0xFFFFFFFFu, none, V.tris); return false; }
    L.slot = raster ? pixel : lt * CLRT_TILE_PIXELS + inTile; L.pixel = pixel;
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    if (!klane_begin(L, V, o, d)) { kd_write_contact(out, L.slot, pixel, none, V.tris); return false; }
    return true;
  }
};
struct KdRaysJob {
  const clrt_ray* rays;
  __device__ __forceinline__ bool start(uint32_t i, KLane& L, const KdView& V, clrt_contact* __restrict__ out) const {
    const float4* p = reinterpret_cast<const float4*>(rays + i);
    const float4 ro = __ldcs(p + 1), rd = __ldcs(p + 2);
    KdHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
    L.slot = i; L.pixel = i;
    if ((rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f) || !klane_begin(L, V, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z))) {
      kd_write_contact(out, i, i, none, V.tris); return false;
    }
    return true;
  }
};
__global__ void __launch_bounds__(KD_THREADS, CLRT_KD_MIN_CTAS)
kd_walk_primary_kernel(const clrt_camera cam, const KdView V, clrt_contact* __restrict__ out, uint32_t rank, uint32_t nranks, int raster,
                       uint32_t* __restrict__ counter) {
  KdPrimaryJob job{cam, make_tile_grid(cam.resX, cam.resY), rank, nranks, raster};
  const uint32_t total = stripe_tile_count(job.tg.tiles, rank, nranks) * 8u * 32u;
  kd_run_jobs(job, V, total, counter, out);
}
__global__ void __launch_bounds__(KD_THREADS, CLRT_KD_MIN_CTAS)
kd_walk_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const KdView V, clrt_contact* __restrict__ out, uint32_t* __restrict__ counter) {
  KdRaysJob job{rays};
  kd_run_jobs(job, V, count, counter, out);
}

} // namespace

// ---- manager -----------------------------------------------------------------------------------------------------------------
static int kd_read_u32(clrt_ctx* ctx, const uint32_t* d, uint32_t* out) {
  void* pin = nullptr;
  CLRT_TRY(clrt_ctx_pinned(ctx, 64, &pin));
  CLRT_CUDA(cudaMemcpyAsync(pin, d, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
  *out = *(uint32_t*)pin;
  return 0;
}
static int kd_grow(clrt_ctx* ctx, void** p, size_t* cap, size_t bytes) {
  if (bytes > *cap || !*p) { CLRT_CUDA(cudaStreamSynchronize(ctx->stream)); CLRT_TRY(clrt_grow(p, cap, bytes ? bytes : 16)); }
  return 0;
}
static KdGrid kd_grid_of(const KdState* k, const clrt_accel* a) {
  KdGrid g; g.box = a->scene->header.modelsBoundingBox; g.lox = k->lo[0]; g.loy = k->lo[1]; g.loz = k->lo[2]; g.csx = k->cs[0]; g.csy = k->cs[1]; g.csz = k->cs[2]; g.levels = k->levels;
  return g;
}

// initializeFrame: the cell grid for the scene as it is now (2^L cells per axis, about two cells per triangle) + buffers.
int kd_init_frame(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  if (!a->kd) a->kd = new KdState();
  KdState* k = a->kd;
  k->built = false; a->built = false;
  const uint32_t n = a->scene->n_tris;
  a->n = n;
  int L = 1;
  while (L < 10 && (double)(1ull << (3 * L)) < 2.0 * (double)(n ? n : 1)) L++;
  k->levels = L;
  const clrt_aabb& box = a->scene->header.modelsBoundingBox;
  const float lo[3] = {box.bounds[0].x, box.bounds[0].y, box.bounds[0].z}, hi[3] = {box.bounds[1].x, box.bounds[1].y, box.bounds[1].z};
  float ext[3], longest = 0.0f;
  for (int i = 0; i < 3; i++) { ext[i] = hi[i] - lo[i]; if (!(ext[i] >= 0.0f)) CLRT_FAIL("KD-tree: the scene box is not a box"); if (ext[i] > longest) longest = ext[i]; }
  if (!(longest < 3.0e37f)) CLRT_FAIL("KD-tree: the scene box is unbounded");
  for (int i = 0; i < 3; i++) {
    float e = ext[i];
    const float floor_e = longest > 0.0f ? longest * 1.0e-6f : 1.0e-30f;      // a flat scene still gets cells of non-zero size
    if (e < floor_e) e = floor_e;
    const float pad = e * 1.0e-4f;
    k->lo[i] = lo[i] - pad;
    k->cs[i] = (e + 2.0f * pad) / (float)(1 << L);
  }
  CLRT_TRY(kd_grow(ctx, (void**)&k->d_counters, &k->counters_cap, (size_t)(n ? n : 1) * 4));
  CLRT_TRY(kd_grow(ctx, (void**)&k->d_scan, &k->scan_cap, (size_t)(n ? n : 1) * 4));
  return 0;
}

// construct: count -> scan -> scatter -> sort -> leaves (run heads -> scan -> codes + ranges) -> radix tree with planes.
int kd_construct(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  KdState* k = a->kd;
  if (!k || !k->d_counters) CLRT_FAIL("KD-tree construct: initializeFrame was not called");
  const uint32_t n = a->n;
  if (n != a->scene->n_tris) CLRT_FAIL("KD-tree construct: scene changed since initializeFrame");
  cudaStream_t s = ctx->stream;
  const KdGrid g = kd_grid_of(k, a);
  const bool tm = a->timing;
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[0], s));
  k->pairs = 0; k->leaves = 0;
  if (n) {
    kd_count_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->scene->d_tris, n, g, k->d_counters);
    CLRT_CHECK_LAUNCH();
    CLRT_TRY(scan_inclusive(ctx, k->d_counters, k->d_scan, n));
    CLRT_TRY(kd_read_u32(ctx, k->d_scan + (n - 1), &k->pairs));
    if (k->pairs >= (1u << 30)) CLRT_FAIL("KD-tree: %u (cell, triangle) pairs", k->pairs);
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_pairs, &k->pairs_capacity, (size_t)k->pairs * 8));
    kd_scatter_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->scene->d_tris, n, g, k->d_scan, k->d_counters, k->pairs, (uint2*)k->d_pairs);
    CLRT_CHECK_LAUNCH();
    CLRT_TRY(sort_pairs_radix(ctx, k->d_pairs, k->pairs, 3 * k->levels));
    // leaves = runs of equal codes
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_counters, &k->counters_cap, (size_t)k->pairs * 4));
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_scan, &k->scan_cap, (size_t)k->pairs * 4));
    kd_heads_kernel<<<(k->pairs + 255) / 256, 256, 0, s>>>((const uint2*)k->d_pairs, k->pairs, k->d_counters);
    CLRT_CHECK_LAUNCH();
    CLRT_TRY(scan_inclusive(ctx, k->d_counters, k->d_scan, k->pairs));
    CLRT_TRY(kd_read_u32(ctx, k->d_scan + (k->pairs - 1), &k->leaves));
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_codes, &k->codes_cap, (size_t)k->leaves * 4));
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_ranges, &k->ranges_cap, (size_t)k->leaves * 8));
    CLRT_TRY(kd_grow(ctx, (void**)&k->d_nodes, &k->nodes_cap, (size_t)k->leaves * sizeof(KdNode)));
    kd_leaves_kernel<<<(k->pairs + 255) / 256, 256, 0, s>>>((const uint2*)k->d_pairs, k->pairs, k->d_scan, k->d_codes, (uint2*)k->d_ranges);
    CLRT_CHECK_LAUNCH();
    if (k->leaves > 1) {
      kd_emit_kernel<<<(k->leaves + 255) / 256, 256, 0, s>>>(k->d_codes, k->leaves, g, k->d_nodes);
      CLRT_CHECK_LAUNCH();
    }
  }
  if (tm) {
    CLRT_CUDA(cudaEventRecord(a->ev[1], s));
    CLRT_CUDA(cudaEventSynchronize(a->ev[1]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_TOTAL], a->ev[0], a->ev[1]));
  }
  k->built = true; a->built = true;
  return 0;
}

template <auto kernel>
static int kd_grid_size(clrt_ctx* ctx, uint32_t packets, unsigned* grid) {
  static int per_sm = 0;
  if (!per_sm) { CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, KD_THREADS, 0)); if (per_sm < 1) per_sm = 1; }
  *grid = min((unsigned)(ctx->sm_count * per_sm), (packets + 3) / 4);
  if (!*grid) *grid = 1;
  return 0;
}

int kd_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks) {
  clrt_ctx* ctx = a->ctx; KdState* k = a->kd;
  if (!k || !k->built) CLRT_FAIL("generateContacts: the KD-tree was not constructed");
  const bool raster = (nranks == 0);
  if (raster) { rank = 0; nranks = 1; }
  const TileGrid tg = make_tile_grid(cam.resX, cam.resY);
  const uint32_t packets = stripe_tile_count(tg.tiles, rank, nranks) * 8u;
  if (!packets) return 0;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  unsigned grid = 0;
#if CLRT_KD_REFILL
  {
    const KdView V{kd_grid_of(k, a), k->d_nodes, k->leaves, (const uint2*)k->d_ranges, (const uint2*)k->d_pairs, a->scene->d_tris, a->d_tile_counter + 13};
    CLRT_TRY(kd_grid_size<kd_walk_primary_kernel>(ctx, packets, &grid));
    kd_walk_primary_kernel<<<grid, KD_THREADS, 0, ctx->stream>>>(cam, V, out, rank, nranks, raster, a->d_tile_counter);
    CLRT_CHECK_LAUNCH();
  }
#else
  CLRT_TRY(kd_grid_size<kd_trace_primary_kernel>(ctx, packets, &grid));
  kd_trace_primary_kernel<<<grid, KD_THREADS, 0, ctx->stream>>>(cam, kd_grid_of(k, a), k->d_nodes, k->leaves, (const uint2*)k->d_ranges,
      (const uint2*)k->d_pairs, a->scene->d_tris, out, rank, nranks, raster, a->d_tile_counter, a->d_tile_counter + 13);
#endif
  CLRT_CHECK_LAUNCH();
  return 0;
}

int kd_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count) {
  clrt_ctx* ctx = a->ctx; KdState* k = a->kd;
  if (!k || !k->built) CLRT_FAIL("generateContacts: the KD-tree was not constructed");
  if (!count) return 0;
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  unsigned grid = 0;
#if CLRT_KD_REFILL
  {
    const KdView V{kd_grid_of(k, a), k->d_nodes, k->leaves, (const uint2*)k->d_ranges, (const uint2*)k->d_pairs, a->scene->d_tris, a->d_tile_counter + 13};
    CLRT_TRY(kd_grid_size<kd_walk_rays_kernel>(ctx, packets, &grid));
    kd_walk_rays_kernel<<<grid, KD_THREADS, 0, ctx->stream>>>(rays, count, V, out, a->d_tile_counter);
    CLRT_CHECK_LAUNCH();
  }
#else
  CLRT_TRY(kd_grid_size<kd_trace_rays_kernel>(ctx, packets, &grid));
  kd_trace_rays_kernel<<<grid, KD_THREADS, 0, ctx->stream>>>(rays, count, kd_grid_of(k, a), k->d_nodes, k->leaves, (const uint2*)k->d_ranges,
      (const uint2*)k->d_pairs, a->scene->d_tris, out, a->d_tile_counter, a->d_tile_counter + 13);
#endif
  CLRT_CHECK_LAUNCH();
  return 0;
}

void kd_destroy(clrt_accel* a) {
  KdState* k = a->kd;
  if (!k) return;
  void* ptrs[] = {k->d_counters, k->d_scan, k->d_pairs, k->d_codes, k->d_ranges, k->d_nodes};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete k;
  a->kd = nullptr;
}

// Inspection (tests): {levels, pairs, leaves}; grid origin + cell size.
extern "C" int clrt_kd_info(clrt_accel* a, uint32_t* out3, float* out_lo_cs6) {
  if (!a || a->kind != ACCEL_KD || !a->kd) CLRT_FAIL("clrt_kd_info: not a KD-tree");
  if (out3) { out3[0] = (uint32_t)a->kd->levels; out3[1] = a->kd->pairs; out3[2] = a->kd->leaves; }
  if (out_lo_cs6) for (int i = 0; i < 3; i++) { out_lo_cs6[i] = a->kd->lo[i]; out_lo_cs6[3 + i] = a->kd->cs[i]; }
  return 0;
}
