// trace_tree.cu — the separate traversal hierarchy ("trace tree"; ours, the reference has none).
//
// Why: the reference's Morton code is c/(max-min) WITHOUT subtracting min (BVH.h:112-114), so every
// axis on which the scene box reaches below zero collapses to cell 0 for all negative centroids, and equal codes become
// right-leaning chains (BVH.h:153-154,212-233, Q8).  On ClumsyDragon 20 % of the triangles sit in 174 runs of 32..211
// equal keys whose boxes are slabs through the model: a ray that enters one tests every member.  The reference-layout
// tree (bvh.cu) stays bit-exact — it is what construct() is specified to produce — but it is a poor thing to traverse.
//
// What: when the reference tree has long runs or collapsed axes, construct() also builds a plain LBVH over the SAME
// triangles from a proper 30-bit Morton code ((c-min)/(max-min), 10 bits per axis) with the usual index tie-break for
// equal codes (Karras 2012, section 4), so there are no chains.  The closest hit does not depend on the hierarchy; exact
// ties are resolved by the triangle's position in the REFERENCE's sorted order (kept in v2.w), i.e. the reference's
// winner (trace.cu).  Traversal kernels take it through the same TraceNode / clrt_triangle layouts.
//
// Kernels: tt_morton_kernel (48 B in, 8 B out per triangle) -> onesweep sort (sort.cu) -> tt_emit_kernel (Karras
// ranges + gather of the triangles in tree order; both child boxes of a node by range queries on the box pyramid of the
// leaf boxes in this tree's order, boxpyr.cuh — no bottom-up pass; a child's box is stored INSIDE its parent's TraceNode: 64 B per node).
#include "common.cuh"
#include "rtmath.cuh"
#include "boxpyr.cuh"

namespace {

__device__ __forceinline__ uint32_t tt_expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t tt_quantise(float c, float lo, float inv_len) {
  const float q = (c - lo) * inv_len * 1024.0f;
  return (uint32_t)fminf(fmaxf(q, 0.0f), 1023.0f);
}

__global__ void __launch_bounds__(256)
tt_morton_kernel(const clrt_triangle* __restrict__ tris, uint32_t n, clrt_aabb bb, uint2* __restrict__ pairs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4* t = reinterpret_cast<const float4*>(tris + i);
  const float4 a = __ldg(t), b = __ldg(t + 1), c = __ldg(t + 2);
  const float lx = bb.bounds[1].x - bb.bounds[0].x, ly = bb.bounds[1].y - bb.bounds[0].y, lz = bb.bounds[1].z - bb.bounds[0].z;
  const uint32_t qx = tt_quantise((a.x + b.x + c.x) * (1.0f / 3.0f), bb.bounds[0].x, lx > 0.0f ? 1.0f / lx : 0.0f);
  const uint32_t qy = tt_quantise((a.y + b.y + c.y) * (1.0f / 3.0f), bb.bounds[0].y, ly > 0.0f ? 1.0f / ly : 0.0f);
  const uint32_t qz = tt_quantise((a.z + b.z + c.z) * (1.0f / 3.0f), bb.bounds[0].z, lz > 0.0f ? 1.0f / lz : 0.0f);
  pairs[i] = make_uint2(tt_expand_bits(qx) * 4 + tt_expand_bits(qy) * 2 + tt_expand_bits(qz), i);
}

// delta(i, j): length of the common prefix of the (code, position) keys; -1 outside the array (Karras 2012).
__device__ __forceinline__ int tt_delta(const uint2* __restrict__ list, int n, uint32_t ki, int i, int j) {
  if (j < 0 || j >= n) return -1;
  const uint32_t kj = __ldg(&list[j].x);
  return ki != kj ? __clz(ki ^ kj) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

// triangle -> position in the REFERENCE tree's sorted order (the tie-break key every traversal-side triangle carries).  Only this tree
// needs the table (the reference tree's own triangle copy gets the position directly), so the scatter is paid only when it is built.
__global__ void __launch_bounds__(256)
tt_refpos_kernel(const uint2* __restrict__ ref_list, uint32_t n, uint32_t* __restrict__ refpos) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p < n) refpos[__ldg(&ref_list[p].y)] = p;
}

// Leaf boxes in this tree's order for the pyramid: the box morton_leaves_kernel stored in the leaf's reference node.
struct TtLeafBox {
  const uint2* list; const clrt_bvh_node* ref_nodes;
  __device__ __forceinline__ Box6 operator()(uint32_t p) const {
    const float4* lb = reinterpret_cast<const float4*>(ref_nodes + __ldg(&list[p].y)) + 1;
    const float4 lo = __ldg(lb), hi = __ldg(lb + 1);
    Box6 b; b.lx = lo.x; b.ly = lo.y; b.lz = lo.z; b.hx = hi.x; b.hy = hi.y; b.hz = hi.z;
    return b;
  }
};

// One thread per sorted position p: gathers the triangle into tree order; threads p < n-1 also emit inner node p — children, and both
// child boxes as range queries on the pyramid.
__global__ void __launch_bounds__(256)
tt_emit_kernel(const uint2* __restrict__ list, uint32_t n_, const clrt_triangle* __restrict__ tris,
               const uint32_t* __restrict__ refpos, clrt_triangle* __restrict__ tris_out, TraceNode* __restrict__ tnodes, const PyrView pv) {
  const int n = (int)n_;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  {
    const uint32_t leaf = __ldg(&list[i].y);
    const float4* src = reinterpret_cast<const float4*>(tris + leaf);
    float4 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
    b.w = __uint_as_float(leaf);
    c.w = __uint_as_float(__ldg(refpos + leaf));
    float4* dst = reinterpret_cast<float4*>(tris_out + i);
    dst[0] = a; dst[1] = b; dst[2] = c;
  }
  if (i + 1 >= n) return;
  const uint32_t ki = __ldg(&list[i].x);
  const int d = tt_delta(list, n, ki, i, i + 1) - tt_delta(list, n, ki, i, i - 1) >= 0 ? 1 : -1;
  const int dmin = tt_delta(list, n, ki, i, i - d);
  int lmax = 2;
  while (tt_delta(list, n, ki, i, i + lmax * d) > dmin) lmax <<= 1;
  int l = 0;
  for (int t = lmax >> 1; t >= 1; t >>= 1)
    if (tt_delta(list, n, ki, i, i + (l + t) * d) > dmin) l += t;
  const int j = i + l * d;
  const int dnode = tt_delta(list, n, ki, i, j);
  int s = 0;
  for (int t = (l + 1) >> 1; ; t = (t + 1) >> 1) {
    if (tt_delta(list, n, ki, i, i + (s + t) * d) > dnode) s += t;
    if (t <= 1) break;
  }
  const int gamma = i + s * d + min(d, 0);
  const int first = min(i, j), last = max(i, j);
  const bool aLeaf = first == gamma, bLeaf = last == gamma + 1;
  const uint32_t cA = aLeaf ? (uint32_t)gamma : (uint32_t)gamma + n_;
  const uint32_t cB = bLeaf ? (uint32_t)gamma + 1u : (uint32_t)gamma + 1u + n_;
  const Box6 A = aLeaf ? pyr_load(pv, 0, (uint32_t)gamma) : pyr_query(pv, (uint32_t)first, (uint32_t)gamma);
  const Box6 B = bLeaf ? pyr_load(pv, 0, (uint32_t)gamma + 1u) : pyr_query(pv, (uint32_t)gamma + 1u, (uint32_t)last);
  float4* tn = reinterpret_cast<float4*>(tnodes + i);
  __stcg(tn,     make_float4(A.lx, A.ly, A.lz, A.hx));
  __stcg(tn + 1, make_float4(A.hy, A.hz, B.lx, B.ly));
  __stcg(tn + 2, make_float4(B.lz, B.hx, B.hy, B.hz));
  reinterpret_cast<uint4*>(tn)[3] = make_uint4(cA, cB, 0u, 0u);                // no runs in this tree
}

#if CLRT_WIDE_Q16
// The quantisation grid: the root's box (union of its two child boxes in tnodes[0]) enlarged by a sliver, 65535 steps per axis.
__global__ void wide_frame_kernel(const TraceNode* __restrict__ tnodes, QFrame* __restrict__ frame) {
  const float4* p = reinterpret_cast<const float4*>(tnodes);
  const float4 q0 = p[0], q1 = p[1], q2 = p[2];
  const float lo[3] = {fminf(q0.x, q1.z), fminf(q0.y, q1.w), fminf(q0.z, q2.x)}, hi[3] = {fmaxf(q0.w, q2.y), fmaxf(q1.x, q2.z), fmaxf(q1.y, q2.w)};
  float flo[3], step[3];
  for (int a = 0; a < 3; a++) {
    const float pad = (hi[a] - lo[a]) * (1.0f / 4096.0f) + (fabsf(lo[a]) + fabsf(hi[a])) * (1.0f / 1048576.0f) + 1.0e-30f;
    flo[a] = lo[a] - pad;
    step[a] = ((hi[a] + pad) - flo[a]) * (1.0f / 65535.0f) * 1.000002f;
  }
  frame->lox = flo[0]; frame->loy = flo[1]; frame->loz = flo[2]; frame->pad0 = 0.0f;
  frame->stepx = step[0]; frame->stepy = step[1]; frame->stepz = step[2]; frame->pad1 = 0.0f;
}
// One bound pair on one axis -> 16-bit grid positions whose decoding lo + q * step (one fused rounding) contains [lo, hi].
__device__ __forceinline__ void q16_pair(float flo, float step, float lo, float hi, bool valid, uint32_t& ql, uint32_t& qh) {
  if (!valid) { ql = 65535u; qh = 0u; return; }
  const float inv = 1.0f / step;
  float l = fminf(fmaxf(floorf((lo - flo) * inv), 0.0f), 65535.0f), h = fminf(fmaxf(ceilf((hi - flo) * inv), 0.0f), 65535.0f);
  while (l > 0.0f && fmaf(l, step, flo) > lo) l -= 1.0f;
  while (h < 65535.0f && fmaf(h, step, flo) < hi) h += 1.0f;
  ql = (uint32_t)l; qh = (uint32_t)h;
}
#endif

// Collapse of the binary trace structure into 4-wide nodes.  One thread per binary inner node i: wide node i starts from i's two
// children and twice replaces the INNER slot with the largest surface area by that child's two children (the usual greedy collapse:
// the child most likely to be entered anyway is opened).  Round 2a opened both children unconditionally — the grandchildren — so a
// leaf child left a slot empty and nodes near the leaves, where a ray spends most of its visits, held two or three boxes.  3 x 64 B
// in, 64 B out; no ordering between nodes is needed because EVERY binary node gets a wide node — the ones not reachable through
// wide links are never visited.  Slots are named registers moved by selects: a running slot index put the node in local memory.
#ifndef CLRT_WIDEN_GREEDY
#define CLRT_WIDEN_GREEDY 1
#endif
struct WSlot { float4 lo, hi; uint32_t c; };                       // .xyz used; c: leaf (< n), inner (n + index) or 0xFFFFFFFF = empty
__device__ __forceinline__ WSlot wslot_empty() {
  const float qnan = __int_as_float(0x7fc00000);                   // empty slot: a NaN box fails every comparison of the slab test
  WSlot s; s.lo = make_float4(qnan, qnan, qnan, 0.0f); s.hi = s.lo; s.c = 0xFFFFFFFFu; return s;
}
__device__ __forceinline__ float wslot_area(const WSlot& s, uint32_t n) {          // -1 for slots that cannot be opened
  if (s.c < n || s.c == 0xFFFFFFFFu) return -1.0f;
  const float dx = s.hi.x - s.lo.x, dy = s.hi.y - s.lo.y, dz = s.hi.z - s.lo.z;
  return fmaxf(dx * dy + dy * dz + dz * dx, 0.0f);
}
__device__ __forceinline__ void wslot_children(const TraceNode* __restrict__ tnodes, uint32_t n, uint32_t inner, WSlot& x, WSlot& y) {
  const float4* p = reinterpret_cast<const float4*>(tnodes + (inner - n));
  const float4 q0 = __ldg(p), q1 = __ldg(p + 1), q2 = __ldg(p + 2);
  const uint4 q3 = __ldg(reinterpret_cast<const uint4*>(p + 3));
  x.lo = make_float4(q0.x, q0.y, q0.z, 0.0f); x.hi = make_float4(q0.w, q1.x, q1.y, 0.0f); x.c = q3.x;
  y.lo = make_float4(q1.z, q1.w, q2.x, 0.0f); y.hi = make_float4(q2.y, q2.z, q2.w, 0.0f); y.c = q3.y;
}
__device__ __forceinline__ void wslot_swap_if(bool c, WSlot& a, WSlot& b) { if (c) { const WSlot t = a; a = b; b = t; } }

__global__ void __launch_bounds__(128)
widen_kernel(const TraceNode* __restrict__ tnodes, uint32_t n, WideNode* __restrict__ wnodes) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= n) return;
  WSlot s0, s1, s2 = wslot_empty(), s3 = wslot_empty();
  wslot_children(tnodes, n, n + i, s0, s1);
#if CLRT_WIDEN_GREEDY
  // first opening: the larger inner child of the two goes to s1, its children take s1 and s2
  wslot_swap_if(wslot_area(s0, n) > wslot_area(s1, n), s0, s1);
  if (wslot_area(s1, n) >= 0.0f) {
    wslot_children(tnodes, n, s1.c, s1, s2);
    // second opening: the largest inner slot of the three goes to s2, its children take s2 and s3
    wslot_swap_if(wslot_area(s0, n) > wslot_area(s2, n), s0, s2);
    wslot_swap_if(wslot_area(s1, n) > wslot_area(s2, n), s1, s2);
    if (wslot_area(s2, n) >= 0.0f) wslot_children(tnodes, n, s2.c, s2, s3);
  }
#else
  // the grandchildren: both children opened whatever their size
  { const WSlot b = s1;
    if (s0.c >= n) wslot_children(tnodes, n, s0.c, s0, s1); else s1 = wslot_empty();
    s2 = b;
    if (b.c >= n) wslot_children(tnodes, n, b.c, s2, s3); }
#endif
#if CLRT_WIDE_Q16
  const QFrame f = *reinterpret_cast<const QFrame*>(wnodes + n + WIDE_TOP_NODES);
  const uint32_t c[4] = {s0.c, s1.c, s2.c, s3.c};
  const float4 lo[4] = {s0.lo, s1.lo, s2.lo, s3.lo}, hi[4] = {s0.hi, s1.hi, s2.hi, s3.hi};
  uint32_t ql[3][4], qh[3][4];
#pragma unroll
  for (int j = 0; j < 4; j++) {
    const bool valid = c[j] != 0xFFFFFFFFu;
    q16_pair(f.lox, f.stepx, lo[j].x, hi[j].x, valid, ql[0][j], qh[0][j]);
    q16_pair(f.loy, f.stepy, lo[j].y, hi[j].y, valid, ql[1][j], qh[1][j]);
    q16_pair(f.loz, f.stepz, lo[j].z, hi[j].z, valid, ql[2][j], qh[2][j]);
  }
  uint4* o = reinterpret_cast<uint4*>(wnodes + i);
  o[0] = make_uint4(ql[0][0] | (ql[0][1] << 16), ql[0][2] | (ql[0][3] << 16), ql[1][0] | (ql[1][1] << 16), ql[1][2] | (ql[1][3] << 16));
  o[1] = make_uint4(ql[2][0] | (ql[2][1] << 16), ql[2][2] | (ql[2][3] << 16), qh[0][0] | (qh[0][1] << 16), qh[0][2] | (qh[0][3] << 16));
  o[2] = make_uint4(qh[1][0] | (qh[1][1] << 16), qh[1][2] | (qh[1][3] << 16), qh[2][0] | (qh[2][1] << 16), qh[2][2] | (qh[2][3] << 16));
  o[3] = make_uint4(c[0], c[1], c[2], c[3]);
#else
  float4* o = reinterpret_cast<float4*>(wnodes + i);
  o[0] = make_float4(s0.lo.x, s1.lo.x, s2.lo.x, s3.lo.x);
  o[1] = make_float4(s0.lo.y, s1.lo.y, s2.lo.y, s3.lo.y);
  o[2] = make_float4(s0.lo.z, s1.lo.z, s2.lo.z, s3.lo.z);
  o[3] = make_float4(s0.hi.x, s1.hi.x, s2.hi.x, s3.hi.x);
  o[4] = make_float4(s0.hi.y, s1.hi.y, s2.hi.y, s3.hi.y);
  o[5] = make_float4(s0.hi.z, s1.hi.z, s2.hi.z, s3.hi.z);
  reinterpret_cast<uint4*>(o)[6] = make_uint4(s0.c, s1.c, s2.c, s3.c);
  reinterpret_cast<uint4*>(o)[7] = make_uint4(0u, 0u, 0u, 0u);
#endif
}

// Top table (the north star's "staging of the top BVH levels"): the first WIDE_TOP_LEVELS levels of the wide tree copied in
// breadth-first order into WIDE_TOP_NODES consecutive records behind the n per-node records (slot of child j of slot s = 4s+1+j, the
// implicit complete 4-ary layout; slots of absent children stay unused).  Child ids that stay inside the table become
// WIDE_TOP_BIT | slot; everything else keeps its id.  The trace kernels start at WIDE_TOP_BIT | 0 and copy a prefix of the table
// into shared memory with one bulk-async (TMA) copy per CTA (trace.cu).  One CTA, one thread per slot, level by level.
__global__ void __launch_bounds__(WIDE_TOP_NODES <= 1024 ? 512 : 1024)
wide_top_kernel(const WideNode* __restrict__ wnodes, uint32_t n, WideNode* __restrict__ top) {
  __shared__ uint32_t src[WIDE_TOP_NODES];
  for (int s = threadIdx.x; s < WIDE_TOP_NODES; s += blockDim.x) src[s] = 0xFFFFFFFFu;
  __syncthreads();
  if (threadIdx.x == 0) src[0] = n;                          // the root: inner node 0
  __syncthreads();
  int first = 0, count = 1;
  for (int level = 0; level < WIDE_TOP_LEVELS; level++) {
    for (int k = threadIdx.x; k < count; k += blockDim.x) {
      const int s = first + k;
      const uint32_t id = src[s];
      uint4* o = reinterpret_cast<uint4*>(top + s);
      constexpr int CH = sizeof(WideNode) / 16 - (CLRT_WIDE_Q16 ? 1 : 2);     // index of the 16 bytes holding the four child ids
      if (id == 0xFFFFFFFFu) {                                 // unused slot: boxes no ray enters, empty children
#if CLRT_WIDE_Q16
        o[0] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu); o[1] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0u, 0u); o[2] = make_uint4(0u, 0u, 0u, 0u);
#else
        const uint32_t q = 0x7fc00000u;
        for (int w = 0; w < 6; w++) o[w] = make_uint4(q, q, q, q);
        o[7] = make_uint4(0u, 0u, 0u, 0u);
#endif
        o[CH] = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
        continue;
      }
      const uint4* p = reinterpret_cast<const uint4*>(wnodes + (id - n));
      for (int w = 0; w < (int)(sizeof(WideNode) / 16); w++) o[w] = p[w];
      const uint4 ch = p[CH];
      uint32_t c[4] = {ch.x, ch.y, ch.z, ch.w};
      for (int j = 0; j < 4; j++)
        if (level + 1 < WIDE_TOP_LEVELS && c[j] != 0xFFFFFFFFu && c[j] >= n) { src[4 * s + 1 + j] = c[j]; c[j] = WIDE_TOP_BIT | (uint32_t)(4 * s + 1 + j); }
      o[CH] = make_uint4(c[0], c[1], c[2], c[3]);
    }
    __syncthreads();
    first += count; count *= 4;
  }
}

} // namespace

void trace_tree_destroy(clrt_accel* a) {
  if (a->d_tt_nodes) cudaFree(a->d_tt_nodes);
  if (a->d_tt_tris) cudaFree(a->d_tt_tris);
  if (a->d_tt_pairs) cudaFree(a->d_tt_pairs);
  if (a->d_wnodes) cudaFree(a->d_wnodes);
  a->d_wnodes = nullptr;
  a->d_tt_nodes = nullptr; a->d_tt_tris = nullptr; a->d_tt_pairs = nullptr;
}

// Called at the end of bvh_construct (a->max_run is known).  Auto policy: build when the reference tree holds an equal-key
// run of TT_RUN_MIN or more, or when more than 1/16 of the scene box lies below zero on some axis (collapsed axis, Q6).
constexpr uint32_t TT_RUN_MIN = 6;
static int trace_tree_separate(clrt_accel* a);
int trace_tree_build(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  const uint32_t n = a->n;
  a->tt_built = false;
  a->stage_ms[CLRT_STAGE_TRACE_TREE] = 0.0f;
  const clrt_aabb bb = a->scene->header.modelsBoundingBox;
  // an axis counts as collapsed when more than 1/16 of the scene box's extent lies below zero (all of it maps to cell 0)
  auto below = [](float lo, float hi) { return lo < 0.0f && -lo * 16.0f > (hi - lo); };
  const bool collapsed = below(bb.bounds[0].x, bb.bounds[1].x) || below(bb.bounds[0].y, bb.bounds[1].y) || below(bb.bounds[0].z, bb.bounds[1].z);
  const bool want = a->tt_policy == 1 || (a->tt_policy == 0 && (a->max_run >= TT_RUN_MIN || collapsed));
  if (n < 2) return 0;
  cudaStream_t s = ctx->stream;
  const bool tm = a->timing;
  const size_t wide_bytes = ((size_t)n + WIDE_TOP_NODES + 1) * sizeof(WideNode);  // per-node records, the top table, the quantisation grid
  if (wide_bytes > a->wnodes_cap) CLRT_CUDA(cudaStreamSynchronize(s));
  CLRT_TRY(clrt_grow((void**)&a->d_wnodes, &a->wnodes_cap, wide_bytes));
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[5], s));
  if (want) CLRT_TRY(trace_tree_separate(a));
#if CLRT_WIDE_Q16
  wide_frame_kernel<<<1, 1, 0, s>>>(want ? a->d_tt_nodes : a->d_tnodes, reinterpret_cast<QFrame*>(a->d_wnodes + n + WIDE_TOP_NODES));
  CLRT_CHECK_LAUNCH();
#endif
  widen_kernel<<<(n + 127) / 128, 128, 0, s>>>(want ? a->d_tt_nodes : a->d_tnodes, n, a->d_wnodes);
  CLRT_CHECK_LAUNCH();
  wide_top_kernel<<<1, WIDE_TOP_NODES <= 1024 ? 512 : 1024, 0, s>>>(a->d_wnodes, n, a->d_wnodes + n);
  CLRT_CHECK_LAUNCH();
  if (tm) {
    CLRT_CUDA(cudaEventRecord(a->ev[6], s));
    CLRT_CUDA(cudaEventSynchronize(a->ev[6]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_TRACE_TREE], a->ev[5], a->ev[6]));
    a->stage_ms[CLRT_STAGE_TOTAL] += a->stage_ms[CLRT_STAGE_TRACE_TREE];
  }
  a->tt_built = want;
  return 0;
}

static int trace_tree_separate(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  const uint32_t n = a->n;
  const clrt_aabb bb = a->scene->header.modelsBoundingBox;
  cudaStream_t s = ctx->stream;
  if ((size_t)n * sizeof(TraceNode) > a->tt_nodes_cap || (size_t)n * sizeof(clrt_triangle) > a->tt_tris_cap || (size_t)n * 8 > a->tt_pairs_cap ||
      pyr_bytes(n) > a->pyr_cap)
    CLRT_CUDA(cudaStreamSynchronize(s));
  CLRT_TRY(clrt_grow((void**)&a->d_tt_nodes, &a->tt_nodes_cap, (size_t)n * sizeof(TraceNode)));
  CLRT_TRY(clrt_grow((void**)&a->d_tt_tris, &a->tt_tris_cap, (size_t)n * sizeof(clrt_triangle)));
  CLRT_TRY(clrt_grow((void**)&a->d_tt_pairs, &a->tt_pairs_cap, (size_t)n * 8));
  CLRT_TRY(clrt_grow((void**)&a->d_pyr, &a->pyr_cap, pyr_bytes(n)));             // the reference tree's pyramid is done with: reused
  const unsigned grid = (n + 255) / 256;
  tt_morton_kernel<<<grid, 256, 0, s>>>(a->scene->d_tris, n, bb, reinterpret_cast<uint2*>(a->d_tt_pairs));
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(sort_pairs_radix(ctx, a->d_tt_pairs, n, 30));
  tt_refpos_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const uint2*>(a->d_pairs), n, a->d_refpos);
  CLRT_CHECK_LAUNCH();
  const PyrView pv = pyr_view(a->d_pyr, n);
  CLRT_TRY(pyr_build(ctx, pv, TtLeafBox{reinterpret_cast<const uint2*>(a->d_tt_pairs), a->d_nodes}));
  tt_emit_kernel<<<grid, 256, 0, s>>>(reinterpret_cast<const uint2*>(a->d_tt_pairs), n, a->scene->d_tris, a->d_refpos, a->d_tt_tris, a->d_tt_nodes, pv);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_accel_set_trace_tree(clrt_accel* a, int policy) {
  if (!a) CLRT_FAIL("clrt_accel_set_trace_tree: NULL accel");
  if (policy < 0 || policy > 2) CLRT_FAIL("clrt_accel_set_trace_tree: policy %d (0 auto, 1 always, 2 never)", policy);
  a->tt_policy = policy;
  return 0;
}
extern "C" int clrt_accel_has_trace_tree(clrt_accel* a, int* out) {
  if (!a || !out) CLRT_FAIL("clrt_accel_has_trace_tree: NULL argument");
  *out = a->tt_built ? 1 : 0;
  return 0;
}
