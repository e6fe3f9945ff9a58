// bvh.cu — LBVH builder: BVHManager::initializeFrame / construct (R/src/CLRayTracerEngine/BVHManager.cpp:122-228).
//
// Stages (one kernel each, no host synchronisation in between; the reference waits on the host after every launch):
//   morton_leaves_kernel   <- calculateMortonCodes (BVHKernels.cl:46-50 -> BVH.h:83-130)   104 B / triangle
//   onesweep radix sort    <- BitonicSort::sort (BVHManager.cpp:195)                        16 B / pair (algorithmic)
//   box pyramid            <- computeBoundingBoxes (BVHKernels.cl:65-79 -> BVH.h:328-334)   the leaf boxes in sorted order reduced to
//                             aligned power-of-two blocks (boxpyr.cuh): 40 B gathered + 64 B written per leaf
//   karras_emit_kernel     <- buildRadixTree (BVHKernels.cl:55-60 -> BVH.h:138-321)         range + split per inner node, then BOTH
//                             child boxes by range queries on the pyramid: node box + 64 B trace node, no bottom-up pass
// Products: nodes[2N-1] in the reference's 48-byte layout (leaves 0..N-1, inner node i at slot N+i, root = slot N),
// the sorted (morton, leaf) pairs, and — ours — a 64-byte TraceNode per inner node for the traversal kernels.
//
// Parity stance: Q1 FIXED — all N-1 inner nodes are built (the reference's guard skips the
// last one); Q3 — no arrival counters at all: a node's box is a range query (boxpyr.cuh), same bits; Q4 LITERAL — inner boxes start from
// (FLT_MAX.., FLT_MIN..) and that box takes part in the merge, so every max bound is >= FLT_MIN;
// Q6 LITERAL — Morton normalisation is c/(max-min) without subtracting min.
#include "common.cuh"
#include "rtmath.cuh"
#include "boxpyr.cuh"

namespace {


// ---- stage 1: leaf AABB + Morton code + leaf node + (code, leaf) pair --------------------------------------
__global__ void __launch_bounds__(256)
morton_leaves_kernel(const clrt_triangle* __restrict__ tris, const uint32_t* __restrict__ tri_in_mesh, uint32_t n,
                     clrt_aabb sceneBB, clrt_bvh_node* __restrict__ nodes, clrt_uint2* __restrict__ pairs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float4* t = reinterpret_cast<const float4*>(tris + i);
  const float4 a = __ldg(t), b = __ldg(t + 1), c = __ldg(t + 2);
  // calculateTriangleAABB (AABB.h:61-79): MIN3(MIN3(v1,v2),v3) component-wise with the compare-select macros
  float mnx = rmin(rmin(a.x, b.x), c.x), mny = rmin(rmin(a.y, b.y), c.y), mnz = rmin(rmin(a.z, b.z), c.z);
  float mxx = rmax(rmax(a.x, b.x), c.x), mxy = rmax(rmax(a.y, b.y), c.y), mxz = rmax(rmax(a.z, b.z), c.z);
  const float xFlat = (float)(uint32_t)((mxx - mnx) < RT_FLT_EPSILON);
  const float yFlat = (float)(uint32_t)((mxy - mny) < RT_FLT_EPSILON);
  const float zFlat = (float)(uint32_t)((mxz - mnz) < RT_FLT_EPSILON);
  mxx += xFlat * RT_FLT_EPSILON; mxy += yFlat * RT_FLT_EPSILON; mxz += zFlat * RT_FLT_EPSILON;
  mnx -= xFlat * RT_FLT_EPSILON; mny -= yFlat * RT_FLT_EPSILON; mnz -= zFlat * RT_FLT_EPSILON;
  // centroid (BVH.h:101-104) and normalisation to the scene box (:112-114)
  const uint32_t code = reference_morton(a, b, c, sceneBB);
  // leaf node: data = {parent (set by the emit stage), triangle-in-submesh, submesh, model}; type LEAF (1.0f) in bounds[0].w
  uint4* o = reinterpret_cast<uint4*>(nodes + i);
  o[0] = make_uint4(0xFFFFFFFFu, __ldg(tri_in_mesh + i), __float_as_uint(b.w), __float_as_uint(c.w));
  reinterpret_cast<float4*>(o)[1] = make_float4(mnx, mny, mnz, 1.0f);
  reinterpret_cast<float4*>(o)[2] = make_float4(mxx, mxy, mxz, 0.0f);
  reinterpret_cast<uint2*>(pairs)[i] = make_uint2(code, i);
}

// Sorted codes for the searches of the emit stage: the CTA's own neighbourhood [lo, hi) sits in shared memory (most nodes are small: their
// dependent probes — ~40 L2 round trips per thread otherwise — stay on the SM), everything else comes from the sorted (code, leaf) pairs.
#ifndef CLRT_EMIT_KEY_WINDOW
#define CLRT_EMIT_KEY_WINDOW 0       /* positions staged on either side of the CTA's 256 (measured: no effect, profiles/r02_experiments.md 5b); 0: no staging */
#endif
struct KeyList { const uint2* list; const uint32_t* s; int lo, hi; };
__device__ __forceinline__ uint32_t key_at(const KeyList& k, int i) {
  if (CLRT_EMIT_KEY_WINDOW && i >= k.lo && i < k.hi) return k.s[i - k.lo];
  return k.list[i].x;
}

// ---- stage 3: Karras radix tree with the reference's duplicate-key chains ------------------------------------
// findSplit (BVH.h:138-182)
__device__ __forceinline__ int find_split(const KeyList& list, int first, int last) {
  const uint32_t firstCode = key_at(list, first), lastCode = key_at(list, last);
  if (firstCode == lastCode) return first;
  const int commonPrefix = __clz(firstCode ^ lastCode);
  int split = first, step = last - first;
  do {
    step = (step + 1) >> 1;
    const int newSplit = split + step;
    if (newSplit < last) {
      const int splitPrefix = __clz(firstCode ^ key_at(list, newSplit));
      if (splitPrefix > commonPrefix) split = newSplit;
    }
  } while (step > 1);
  return split;
}
// determineRange (BVH.h:190-282)
__device__ __forceinline__ int2 determine_range(const KeyList& list, int index, uint32_t size) {
  const int lso = (int)size - 1;
  if (index == 0) return make_int2(0, lso);
  const int initialindex = index;
  const uint32_t minone = key_at(list, index - 1), precis = key_at(list, index), pluone = key_at(list, index + 1);
  int dir, d_min;
  if (minone == precis && pluone == precis) {
    // inside a run of equal codes: the node covers [index, end of run] (chain)
    while (index > 0 && index < lso) {
      index += 1;
      if (index >= lso || key_at(list, index) != key_at(list, index + 1)) break;
    }
    return make_int2(initialindex, index);
  }
  const int lx = __clz(precis ^ minone), ly = __clz(precis ^ pluone);
  if (lx > ly) { dir = -1; d_min = ly; } else { dir = 1; d_min = lx; }
  int l_max = 2;
  int testindex = index + l_max * dir;
  while ((testindex <= lso && testindex >= 0) ? (__clz(precis ^ key_at(list, testindex)) > d_min) : false) {
    l_max <<= 1;
    testindex = index + l_max * dir;
  }
  int l = 0;
  for (int div = 2; l_max / div >= 1; div <<= 1) {
    const int t = l_max / div;
    const int newTest = index + (l + t) * dir;
    if (newTest <= lso && newTest >= 0) {
      const int splitPrefix = __clz(precis ^ key_at(list, newTest));
      if (splitPrefix > d_min) l = l + t;
    }
  }
  const int other = index + l * dir;
  return make_int2(min(index, other), max(index, other));
}

// constructNode (BVH.h:291-321).  One thread per inner node; all N-1 are built (Q1 FIXED).
// Writes: own children (data[1], data[2]; data[3] = 0), parent links of both children, root parent = UINT_MAX.
// The inner box is NOT written here: refit merges (childA, childB, init box) and writes the result, which is what
// the reference's fill (BVH.h:309-310) followed by merge3 (:333) leaves in memory.
// Also here (ours): the traversal-side products that depend only on the sorted order —
//   tris_sorted[p] = triangle of the leaf at sorted position p (v1.w := original triangle id), so that leaves that are
//   neighbours in the tree are neighbours in memory, and TraceNode child ids (leaf -> sorted position p, inner -> N + i).
// Leaf ids in sorted-position space make "smaller id" == "visited earlier by the reference's A-before-B traversal",
// which is the tie-break the FAST traversal uses to pick the reference's winner on exact ties.
#ifndef CLRT_EMIT_DEBUG
#define CLRT_EMIT_DEBUG 0           /* experiments only (wrong results): bit 0 skip the triangle copy, bit 1 skip the boxes, bit 2 skip the parent links, bit 3 skip the searches */
#endif
#ifndef CLRT_EMIT_MIN_CTAS
#define CLRT_EMIT_MIN_CTAS 1
#endif
__global__ void __launch_bounds__(256, CLRT_EMIT_MIN_CTAS)
karras_emit_kernel(const clrt_uint2* __restrict__ pairs_, uint32_t n, clrt_bvh_node* __restrict__ nodes,
                   const clrt_triangle* __restrict__ tris, clrt_triangle* __restrict__ tris_sorted,
                   TraceNode* __restrict__ tnodes, uint32_t* __restrict__ max_run, const PyrView pv) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  const uint2* list = reinterpret_cast<const uint2*>(pairs_);
  __shared__ uint32_t s_keys[256 + 2 * CLRT_EMIT_KEY_WINDOW + 1];
  KeyList keys;
  keys.list = list; keys.s = s_keys;
  keys.lo = max((int)(blockIdx.x * blockDim.x) - CLRT_EMIT_KEY_WINDOW, 0);
  keys.hi = (int)min((size_t)blockIdx.x * blockDim.x + 256 + CLRT_EMIT_KEY_WINDOW, (size_t)n);
  if (CLRT_EMIT_KEY_WINDOW) {
    for (int i = keys.lo + (int)threadIdx.x; i < keys.hi; i += 256) s_keys[i - keys.lo] = __ldg(&list[i].x);
    __syncthreads();
  }
  if (idx >= n) return;
  if (!(CLRT_EMIT_DEBUG & 1)) {
    const uint32_t leaf = list[idx].y;
    const float4* src = reinterpret_cast<const float4*>(tris + leaf);
    float4 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
    b.w = __uint_as_float(leaf);
    c.w = __uint_as_float(idx);                 // position in the reference's sorted order: the tie-break key of the traversal
    float4* dst = reinterpret_cast<float4*>(tris_sorted + idx);
    dst[0] = a; dst[1] = b; dst[2] = c;
  }
  if (idx + 1 >= n) return;
  const int2 range = (CLRT_EMIT_DEBUG & 8) ? make_int2((int)idx, (int)idx + 1) : determine_range(keys, (int)idx, n);
  const int split = (CLRT_EMIT_DEBUG & 8) ? (int)idx : find_split(keys, range.x, range.y);
  const uint32_t self = idx + n;
  const bool aIsLeaf = (range.x == split);
  const bool bIsLeaf = (range.y == split + 1);
  const uint32_t child_A = aIsLeaf ? list[split].y : (uint32_t)split + n;
  const uint32_t child_B = bIsLeaf ? list[split + 1].y : (uint32_t)split + 1 + n;
  uint32_t* d = nodes[self].data;
  if (idx == 0) d[0] = 0xFFFFFFFFu;
  d[1] = child_A; d[2] = child_B; d[3] = 0u;
  if (!(CLRT_EMIT_DEBUG & 4)) {
    nodes[child_A].data[0] = self;
    nodes[child_B].data[0] = self;
  }
  // Equal-key runs (the reference's right-leaning chains, Q8): an inner child whose whole range carries one key is a
  // chain of its leaves in sorted order.  The traversal may treat it as ONE multi-triangle leaf [start, start+count):
  // pad0 = start of child A's run, pad1 = countA | countB << 16 (0 = not a run; child B's run starts at split+1).
  uint32_t cntA = 0, cntB = 0;
  if (!aIsLeaf && key_at(keys, range.x) == key_at(keys, split)) { cntA = (uint32_t)(split - range.x + 1); if (cntA > 0xFFFFu) cntA = 0; }
  if (!bIsLeaf && key_at(keys, split + 1) == key_at(keys, range.y)) { cntB = (uint32_t)(range.y - split); if (cntB > 0xFFFFu) cntB = 0; }
  if (max(cntA, cntB) >= 2) atomicMax(max_run, max(cntA, cntB));
  // Boxes (computeBoundingBoxes, BVH.h:328-334: merge3(childA, childB, self) with self = (FLT_MAX.., INNER = 0.0f), (FLT_MIN.., 0.0f)):
  // each child's box is the union of the leaf boxes of its range — a leaf's own box, or the clamped range query of an inner child.
  if (CLRT_EMIT_DEBUG & 2) return;
  const Box6 A = aIsLeaf ? pyr_load(pv, 0, (uint32_t)split) : box_clamped(pyr_query(pv, (uint32_t)range.x, (uint32_t)split));
  const Box6 B = bIsLeaf ? pyr_load(pv, 0, (uint32_t)split + 1u) : box_clamped(pyr_query(pv, (uint32_t)split + 1u, (uint32_t)range.y));
  const Box6 U = box_clamped(box_join(A, B));
  // type field: leaves carry LEAF = 1.0f in bounds[0].w and 0.0f in bounds[1].w, so merge3 leaves (+0.0f, +0.0f) in every inner node
  float4* po = reinterpret_cast<float4*>(nodes + self) + 1;
  __stcg(po, make_float4(U.lx, U.ly, U.lz, 0.0f)); __stcg(po + 1, make_float4(U.hx, U.hy, U.hz, 0.0f));
  float4* tn = reinterpret_cast<float4*>(tnodes + idx);
  __stcg(tn,     make_float4(A.lx, A.ly, A.lz, A.hx));
  __stcg(tn + 1, make_float4(A.hy, A.hz, B.lx, B.ly));
  __stcg(tn + 2, make_float4(B.lz, B.hx, B.hy, B.hz));
  reinterpret_cast<uint4*>(tn)[3] = make_uint4(aIsLeaf ? (uint32_t)split : (uint32_t)split + n,
                                               bIsLeaf ? (uint32_t)split + 1u : (uint32_t)split + 1u + n,
                                               (uint32_t)range.x, cntA | (cntB << 16));
}

// Leaf boxes in sorted order for the pyramid: the box morton_leaves_kernel stored in the leaf's node.
struct RefLeafBox {
  const uint2* list; const clrt_bvh_node* nodes;
  __device__ __forceinline__ Box6 operator()(uint32_t p) const {
    const float4* lb = reinterpret_cast<const float4*>(nodes + __ldg(&list[p].y)) + 1;
    const float4 lo = __ldg(lb), hi = __ldg(lb + 1);
    Box6 b; b.lx = lo.x; b.ly = lo.y; b.lz = lo.z; b.hx = hi.x; b.hy = hi.y; b.hz = hi.z;
    return b;
  }
};

} // namespace

size_t pyr_bytes(uint32_t n) { return 2 * (size_t)pow2ceil_u32(n ? n : 1) * 32; }
PyrView pyr_view(void* mem, uint32_t n) {
  PyrView pv;
  pv.e = reinterpret_cast<float4*>(mem); pv.n = n; pv.n2 = pow2ceil_u32(n ? n : 1);
  pv.levels = 1;
  while (n > (1ull << (pv.levels - 1))) pv.levels++;
  return pv;
}
// ---- manager ---------------------------------------------------------------------------------------------------
int bvh_init_frame(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  const uint32_t n = a->scene->n_tris;
  a->n = n;
  a->built = false;
  a->tt_built = false;
  a->pairs_items = pow2ceil_u32(n ? n : 1);                  // the reference pads to 2^k for its bitonic sort (BVHManager.cpp:129-131)
  const size_t nodes_bytes = (size_t)(n ? 2 * (size_t)n - 1 : 1) * sizeof(clrt_bvh_node);
  if (nodes_bytes > a->nodes_cap || (size_t)a->pairs_items * 8 > a->pairs_cap ||
      (size_t)n * 4 > a->refpos_cap || (size_t)n * sizeof(TraceNode) > a->tnodes_cap || (size_t)n * sizeof(clrt_triangle) > a->tris_sorted_cap ||
      pyr_bytes(n) > a->pyr_cap)
    CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
  CLRT_TRY(clrt_grow((void**)&a->d_nodes, &a->nodes_cap, nodes_bytes));
  CLRT_TRY(clrt_grow((void**)&a->d_pairs, &a->pairs_cap, (size_t)a->pairs_items * 8));
  CLRT_TRY(clrt_grow((void**)&a->d_refpos, &a->refpos_cap, (size_t)(n ? n : 1) * 4));
  CLRT_TRY(clrt_grow((void**)&a->d_tnodes, &a->tnodes_cap, (size_t)(n ? n : 1) * sizeof(TraceNode)));
  CLRT_TRY(clrt_grow((void**)&a->d_tris_sorted, &a->tris_sorted_cap, (size_t)(n ? n : 1) * sizeof(clrt_triangle)));
  CLRT_TRY(clrt_grow((void**)&a->d_pyr, &a->pyr_cap, pyr_bytes(n)));
  // fill of initializeFrame (BVHManager.cpp:138-147): pairs = (UINT_MAX,UINT_MAX).  The reference also zeroes its arrival counters;
  // this builder has none (boxpyr.cuh).
  CLRT_CUDA(cudaMemsetAsync(a->d_pairs, 0xFF, (size_t)a->pairs_items * 8, ctx->stream));
  return 0;
}

int bvh_construct(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  const uint32_t n = a->n;
  if (!a->d_nodes) CLRT_FAIL("BVH construct: initializeFrame was not called");
  if (n != a->scene->n_tris) CLRT_FAIL("BVH construct: scene changed since initializeFrame");
  cudaStream_t s = ctx->stream;
  const bool tm = a->timing;
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[0], s));
  if (n) {
    morton_leaves_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->scene->d_tris, a->scene->d_triInMesh, n,
                                                        a->scene->header.modelsBoundingBox, a->d_nodes, a->d_pairs);
    CLRT_CHECK_LAUNCH();
  }
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[1], s));
  // Only the N real pairs are sorted: the (UINT_MAX,UINT_MAX) padding is already at the end, which is where the
  // reference's sort over 2^k items leaves it.  Morton keys have 30 bits.
  CLRT_TRY(sort_pairs_radix(ctx, a->d_pairs, n, 30));
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[2], s));
  // stage "refit" = the box pyramid over the leaf boxes in sorted order; stage "emit" = radix tree + every node's boxes from it
  const PyrView pv = pyr_view(a->d_pyr, n);
  if (n > 1) CLRT_TRY(pyr_build(ctx, pv, RefLeafBox{reinterpret_cast<const uint2*>(a->d_pairs), a->d_nodes}));
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[3], s));
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter + 12, 0, 4, s));
  if (n) {
    karras_emit_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->d_pairs, n, a->d_nodes, a->scene->d_tris, a->d_tris_sorted, a->d_tnodes, a->d_tile_counter + 12, pv);
    CLRT_CHECK_LAUNCH();
  }
  if (tm) {
    CLRT_CUDA(cudaEventRecord(a->ev[4], s));
    CLRT_CUDA(cudaEventSynchronize(a->ev[4]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_MORTON], a->ev[0], a->ev[1]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_SORT], a->ev[1], a->ev[2]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_REFIT], a->ev[2], a->ev[3]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_EMIT], a->ev[3], a->ev[4]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_TOTAL], a->ev[0], a->ev[4]));
  }
  // longest equal-key run: selects the traversal variant (4-byte read-back; construct() is synchronous in the reference too)
  {
    void* pin = nullptr;
    CLRT_TRY(clrt_ctx_pinned(ctx, 64, &pin));
    CLRT_CUDA(cudaMemcpyAsync(pin, a->d_tile_counter + 12, 4, cudaMemcpyDeviceToHost, s));
    CLRT_CUDA(cudaStreamSynchronize(s));
    a->max_run = *(uint32_t*)pin;
  }
  CLRT_TRY(trace_tree_build(a));
  a->built = true;
  return 0;
}

extern "C" int clrt_bvh_export_nodes(clrt_accel* a, void* h_nodes) {
  if (!a || !h_nodes) CLRT_FAIL("clrt_bvh_export_nodes: NULL argument");
  if (a->kind != ACCEL_BVH || !a->built) CLRT_FAIL("clrt_bvh_export_nodes: not a constructed BVH");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  const size_t bytes = (size_t)(a->n ? 2 * (size_t)a->n - 1 : 0) * sizeof(clrt_bvh_node);
  if (bytes) CLRT_CUDA(cudaMemcpyAsync(h_nodes, a->d_nodes, bytes, cudaMemcpyDeviceToHost, a->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(a->ctx->stream));
  return 0;
}
extern "C" int clrt_bvh_export_pairs(clrt_accel* a, void* h_pairs) {
  if (!a || !h_pairs) CLRT_FAIL("clrt_bvh_export_pairs: NULL argument");
  if (a->kind != ACCEL_BVH || !a->built) CLRT_FAIL("clrt_bvh_export_pairs: not a constructed BVH");
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  if (a->n) CLRT_CUDA(cudaMemcpyAsync(h_pairs, a->d_pairs, (size_t)a->n * 8, cudaMemcpyDeviceToHost, a->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(a->ctx->stream));
  return 0;
}
