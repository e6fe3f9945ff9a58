// boxpyr.cuh — box pyramid: the bounding box of ANY range of leaves in tree (sorted) order with a handful of independent loads.
//
// Why: in a radix tree over sorted keys every inner node covers a contiguous range of leaves [first, last], so its box is the union of
// the leaf boxes of that range.  The reference (computeBoundingBoxes, BVHKernels.cl:65-79 -> BVH.h:328-334) and rounds 1-2a of this
// library compute it bottom-up: one thread per leaf climbs, an arrival counter per node lets the second arrival continue.  That is a
// chain of (fence + atomic + dependent load) per level — ~2.5 us x tree depth on the critical path, 0.16 ms of a 0.44 ms build at
// 1M triangles with the DRAM 10 % busy.  With a pyramid of aligned power-of-two blocks (level k entry j = box of leaves
// [j 2^k, (j+1) 2^k)) a node's box is a RANGE QUERY: at most two entries per level, their addresses known from (first, last) alone —
// no atomics, no counters, no ordering between nodes, and the emit kernel that finds (first, split, last) writes the boxes itself.
//
// Bit-exactness with the reference's merge order: the reference joins with compare-select macros (rmin / rmax, rtmath.cuh) whose ties
// (+0 / -0) go to the SECOND operand, child A (lower positions) first.  "Ties go to the right operand" is associative, so any
// bracketing that keeps the operands in position order returns the same bits as the tree's own bracketing; the query below joins
// left parts left-to-right and right parts right-to-left and never swaps sides.  The (FLT_MAX.., FLT_MIN..) start box of an inner
// node clamps; clamping at every level equals clamping the result, so a child box = clamp(range union) when the child
// is an inner node and the plain leaf box when it is a leaf.
#pragma once
#include "common.cuh"
#include "rtmath.cuh"

#ifndef CLRT_PYR_LDG
#define CLRT_PYR_LDG 1
#endif

struct Box6 { float lx, ly, lz, hx, hy, hz; };
__device__ __forceinline__ Box6 box_join(const Box6& L, const Box6& R) {      // L = lower sorted positions, R = higher
  Box6 o;
  o.lx = rmin(L.lx, R.lx); o.ly = rmin(L.ly, R.ly); o.lz = rmin(L.lz, R.lz);
  o.hx = rmax(L.hx, R.hx); o.hy = rmax(L.hy, R.hy); o.hz = rmax(L.hz, R.hz);
  return o;
}
__device__ __forceinline__ Box6 box_clamped(const Box6& b) {                  // the (FLT_MAX.., FLT_MIN..) start box takes part (Q4)
  Box6 o;
  o.lx = rmin(b.lx, RT_FLT_MAX); o.ly = rmin(b.ly, RT_FLT_MAX); o.lz = rmin(b.lz, RT_FLT_MAX);
  o.hx = rmax(b.hx, RT_FLT_MIN); o.hy = rmax(b.hy, RT_FLT_MIN); o.hz = rmax(b.hz, RT_FLT_MIN);
  return o;
}

// Entry (k, j) lives at index 2 N2 - (2 N2 >> k) + j with N2 = n rounded up to a power of two (levels padded to powers of two: a
// closed form instead of an offset table; the padding is never written or read).  An entry is two float4 (lo.xyz, hi.xyz).
struct PyrView {
  float4* e;
  uint32_t n, n2;                 // leaves, pow2ceil(n)
  int levels;                     // levels 0 .. levels-1; the last one has a single entry
};
__host__ __device__ __forceinline__ size_t pyr_index(const PyrView& pv, int k, uint32_t j) {
  const size_t two = 2 * (size_t)pv.n2;
  return two - (two >> k) + j;
}
__device__ __forceinline__ Box6 pyr_load(const PyrView& pv, int k, uint32_t j) {
  const float4* p = pv.e + 2 * pyr_index(pv, k, j);
#if CLRT_PYR_LDG
  const float4 lo = __ldg(p), hi = __ldg(p + 1);       // entries are written by EARLIER launches only: the read-only path is safe
#else
  const float4 lo = __ldcg(p), hi = __ldcg(p + 1);
#endif
  Box6 b; b.lx = lo.x; b.ly = lo.y; b.lz = lo.z; b.hx = hi.x; b.hy = hi.y; b.hz = hi.z;
  return b;
}
__device__ __forceinline__ void pyr_store(const PyrView& pv, int k, uint32_t j, const Box6& b) {
  float4* p = pv.e + 2 * pyr_index(pv, k, j);
  __stcg(p, make_float4(b.lx, b.ly, b.lz, 0.0f)); __stcg(p + 1, make_float4(b.hx, b.hy, b.hz, 0.0f));
}
// Union of the leaf boxes of positions l .. r (inclusive, r < n), operands in position order.  The walk over the levels depends only on
// (l, r), never on loaded data, so PYR_BATCH levels are unrolled with their (predicated) loads side by side: a range of up to
// 2^PYR_BATCH leaves costs one round trip to L2, not one per level.
#ifndef CLRT_PYR_BATCH
#define CLRT_PYR_BATCH 1
#endif
constexpr int PYR_BATCH = CLRT_PYR_BATCH;
__device__ __forceinline__ Box6 pyr_query(const PyrView& pv, uint32_t l, uint32_t r) {
  uint32_t lo = l, hi = r + 1u;
  Box6 L, R;
  L.lx = L.ly = L.lz = L.hx = L.hy = L.hz = 0.0f; R = L;
  bool hasL = false, hasR = false;
  for (int k = 0; lo < hi; k += PYR_BATCH) {
    Box6 eL[PYR_BATCH], eR[PYR_BATCH];
    bool uL[PYR_BATCH], uR[PYR_BATCH];
#pragma unroll
    for (int q = 0; q < PYR_BATCH; q++) {
      const bool active = lo < hi;
      uL[q] = active && (lo & 1u); uR[q] = active && (hi & 1u);
      eL[q] = L; eR[q] = L;
      if (uL[q]) eL[q] = pyr_load(pv, k + q, lo);
      if (uR[q]) eR[q] = pyr_load(pv, k + q, hi - 1u);
      lo = (lo + 1u) >> 1; hi >>= 1;                   // odd lo: entry taken, next block; odd hi: entry hi-1 taken
    }
#pragma unroll
    for (int q = 0; q < PYR_BATCH; q++) {
      if (uL[q]) { L = hasL ? box_join(L, eL[q]) : eL[q]; hasL = true; }
      if (uR[q]) { R = hasR ? box_join(eR[q], R) : eR[q]; hasR = true; }
    }
  }
  return hasL ? (hasR ? box_join(L, R) : L) : R;
}

// ---- construction ----------------------------------------------------------------------------------------------------------------
// One CTA per 1024 consecutive entries of level `base`: they come from `leaf(p)` when base == 0 (a gather for the reference tree;
// level 0 is written out too) or from the pyramid itself; the ten levels above are reduced in shared memory and written out.
// Launched once per ten levels: 2 launches up to 2^20 leaves, 3 up to 2^30.
constexpr int PYR_SPAN_LOG = 10, PYR_SPAN = 1 << PYR_SPAN_LOG, PYR_THREADS = 256;

template <class LeafBox, bool FROM_LEAVES>        // FROM_LEAVES: base_level == 0 (a compile-time switch keeps the four gathers of a thread in flight together)
__global__ void __launch_bounds__(PYR_THREADS)
pyr_levels_kernel(const LeafBox leaf, const PyrView pv, const int base_level) {
  __shared__ float4 s_lo[PYR_SPAN], s_hi[PYR_SPAN];
  const uint32_t base = blockIdx.x * (uint32_t)PYR_SPAN;
  const uint32_t count = (uint32_t)(((size_t)pv.n + ((size_t)1 << base_level) - 1) >> base_level);     // entries of the base level
#pragma unroll
  for (int q = 0; q < PYR_SPAN / PYR_THREADS; q++) {
    const uint32_t t = threadIdx.x + q * PYR_THREADS, p = base + t;
    if (p < count) {
      if (FROM_LEAVES) {
        const Box6 b = leaf(p);
        s_lo[t] = make_float4(b.lx, b.ly, b.lz, 0.0f); s_hi[t] = make_float4(b.hx, b.hy, b.hz, 0.0f);
        float4* o = pv.e + 2 * pyr_index(pv, 0, p);
        __stcg(o, s_lo[t]); __stcg(o + 1, s_hi[t]);
      } else {
        const float4* e = pv.e + 2 * pyr_index(pv, base_level, p);
        s_lo[t] = __ldcg(e); s_hi[t] = __ldcg(e + 1);
      }
    }
  }
  __syncthreads();
  const int top = min(PYR_SPAN_LOG, pv.levels - 1 - base_level);
  for (int k = 1; k <= top; k++) {
    // entry t of this CTA at level base+k sits in slot t << k; its right half starts 1 << (k-1) slots further (absent beyond the end: copy)
    for (uint32_t t = threadIdx.x; t < (uint32_t)(PYR_SPAN >> k); t += PYR_THREADS) {
      const uint32_t slot = t << k, right = slot + (1u << (k - 1));
      if (base + slot < count) {
        float4 lo = s_lo[slot], hi = s_hi[slot];
        if (base + right < count) {
          const float4 rl = s_lo[right], rh = s_hi[right];
          lo.x = rmin(lo.x, rl.x); lo.y = rmin(lo.y, rl.y); lo.z = rmin(lo.z, rl.z);
          hi.x = rmax(hi.x, rh.x); hi.y = rmax(hi.y, rh.y); hi.z = rmax(hi.z, rh.z);
          s_lo[slot] = lo; s_hi[slot] = hi;
        }
        float4* o = pv.e + 2 * pyr_index(pv, base_level + k, (base >> k) + t);
        __stcg(o, lo); __stcg(o + 1, hi);
      }
    }
    __syncthreads();
  }
}

size_t pyr_bytes(uint32_t n);
PyrView pyr_view(void* mem, uint32_t n);
template <class LeafBox>
int pyr_build(clrt_ctx* ctx, const PyrView& pv, const LeafBox& leaf) {
  if (pv.n == 0) return 0;
  for (int base = 0; base == 0 || base < pv.levels - 1; base += PYR_SPAN_LOG) {
    const size_t count = ((size_t)pv.n + ((size_t)1 << base) - 1) >> base;
    const unsigned grid = (unsigned)((count + PYR_SPAN - 1) / PYR_SPAN);
    if (base == 0) pyr_levels_kernel<LeafBox, true><<<grid, PYR_THREADS, 0, ctx->stream>>>(leaf, pv, 0);
    else pyr_levels_kernel<LeafBox, false><<<grid, PYR_THREADS, 0, ctx->stream>>>(leaf, pv, base);
    CLRT_CHECK_LAUNCH();
  }
  return 0;
}
