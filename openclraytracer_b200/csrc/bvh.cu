// bvh.cu — LBVH builder: BVHManager::initializeFrame / construct (R/src/CLRayTracerEngine/BVHManager.cpp:122-228).
//
// Stages (one kernel each, no host synchronisation in between; the reference waits on the host after every launch):
//   morton_leaves_kernel   <- calculateMortonCodes (BVHKernels.cl:46-50 -> BVH.h:83-130)   104 B / triangle
//   onesweep radix sort    <- BitonicSort::sort (BVHManager.cpp:195)                        16 B / pair (algorithmic)
//   karras_emit_kernel     <- buildRadixTree (BVHKernels.cl:55-60 -> BVH.h:138-321)         64 B / inner node
//   refit_atomic_kernel    <- computeBoundingBoxes (BVHKernels.cl:65-79 -> BVH.h:328-334)   100 B / inner node (+64 B trace node)
// Products: nodes[2N-1] in the reference's 48-byte layout (leaves 0..N-1, inner node i at slot N+i, root = slot N),
// the sorted (morton, leaf) pairs, and — ours — a 64-byte TraceNode per inner node for the traversal kernels.
//
// Parity stance: Q1 FIXED — all N-1 inner nodes are built (the reference's guard skips the
// last one); Q3 — refit uses the atomic's return value + fences; Q4 LITERAL — inner boxes start from
// (FLT_MAX.., FLT_MIN..) and that box takes part in the merge, so every max bound is >= FLT_MIN;
// Q6 LITERAL — Morton normalisation is c/(max-min) without subtracting min.
#include "common.cuh"
#include "rtmath.cuh"

namespace {

__device__ __forceinline__ float4 ldcg4(const float4* p) { return __ldcg(p); }

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

// ---- stage 3: Karras radix tree with the reference's duplicate-key chains ------------------------------------
// findSplit (BVH.h:138-182)
__device__ __forceinline__ int find_split(const uint2* __restrict__ list, int first, int last) {
  const uint32_t firstCode = list[first].x, lastCode = list[last].x;
  if (firstCode == lastCode) return first;
  const int commonPrefix = __clz(firstCode ^ lastCode);
  int split = first, step = last - first;
  do {
    step = (step + 1) >> 1;
    const int newSplit = split + step;
    if (newSplit < last) {
      const int splitPrefix = __clz(firstCode ^ list[newSplit].x);
      if (splitPrefix > commonPrefix) split = newSplit;
    }
  } while (step > 1);
  return split;
}
// determineRange (BVH.h:190-282)
__device__ __forceinline__ int2 determine_range(const uint2* __restrict__ list, int index, uint32_t size) {
  const int lso = (int)size - 1;
  if (index == 0) return make_int2(0, lso);
  const int initialindex = index;
  const uint32_t minone = list[index - 1].x, precis = list[index].x, pluone = list[index + 1].x;
  int dir, d_min;
  if (minone == precis && pluone == precis) {
    // inside a run of equal codes: the node covers [index, end of run] (chain)
    while (index > 0 && index < lso) {
      index += 1;
      if (index >= lso || list[index].x != list[index + 1].x) break;
    }
    return make_int2(initialindex, index);
  }
  const int lx = __clz(precis ^ minone), ly = __clz(precis ^ pluone);
  if (lx > ly) { dir = -1; d_min = ly; } else { dir = 1; d_min = lx; }
  int l_max = 2;
  int testindex = index + l_max * dir;
  while ((testindex <= lso && testindex >= 0) ? (__clz(precis ^ list[testindex].x) > d_min) : false) {
    l_max <<= 1;
    testindex = index + l_max * dir;
  }
  int l = 0;
  for (int div = 2; l_max / div >= 1; div <<= 1) {
    const int t = l_max / div;
    const int newTest = index + (l + t) * dir;
    if (newTest <= lso && newTest >= 0) {
      const int splitPrefix = __clz(precis ^ list[newTest].x);
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
__global__ void __launch_bounds__(256)
karras_emit_kernel(const clrt_uint2* __restrict__ pairs_, uint32_t n, clrt_bvh_node* __restrict__ nodes,
                   const clrt_triangle* __restrict__ tris, clrt_triangle* __restrict__ tris_sorted,
                   TraceNode* __restrict__ tnodes, uint32_t* __restrict__ max_run, uint32_t* __restrict__ refpos) {
  const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  const uint2* list = reinterpret_cast<const uint2*>(pairs_);
  {
    const uint32_t leaf = list[idx].y;
    const float4* src = reinterpret_cast<const float4*>(tris + leaf);
    float4 a = __ldg(src), b = __ldg(src + 1), c = __ldg(src + 2);
    b.w = __uint_as_float(leaf);
    c.w = __uint_as_float(idx);                 // position in the reference's sorted order: the tie-break key of the traversal
    refpos[leaf] = idx;
    float4* dst = reinterpret_cast<float4*>(tris_sorted + idx);
    dst[0] = a; dst[1] = b; dst[2] = c;
  }
  if (idx + 1 >= n) return;
  const int2 range = determine_range(list, (int)idx, n);
  const int split = find_split(list, range.x, range.y);
  const uint32_t self = idx + n;
  const bool aIsLeaf = (range.x == split);
  const bool bIsLeaf = (range.y == split + 1);
  const uint32_t child_A = aIsLeaf ? list[split].y : (uint32_t)split + n;
  const uint32_t child_B = bIsLeaf ? list[split + 1].y : (uint32_t)split + 1 + n;
  uint32_t* d = nodes[self].data;
  if (idx == 0) d[0] = 0xFFFFFFFFu;
  d[1] = child_A; d[2] = child_B; d[3] = 0u;
  nodes[child_A].data[0] = self;
  nodes[child_B].data[0] = self;
  // Equal-key runs (the reference's right-leaning chains, Q8): an inner child whose whole range carries one key is a
  // chain of its leaves in sorted order.  The traversal may treat it as ONE multi-triangle leaf [start, start+count):
  // pad0 = start of child A's run, pad1 = countA | countB << 16 (0 = not a run; child B's run starts at split+1).
  uint32_t cntA = 0, cntB = 0;
  if (!aIsLeaf && list[range.x].x == list[split].x) { cntA = (uint32_t)(split - range.x + 1); if (cntA > 0xFFFFu) cntA = 0; }
  if (!bIsLeaf && list[split + 1].x == list[range.y].x) { cntB = (uint32_t)(range.y - split); if (cntB > 0xFFFFu) cntB = 0; }
  if (max(cntA, cntB) >= 2) atomicMax(max_run, max(cntA, cntB));
  reinterpret_cast<uint4*>(tnodes + idx)[3] = make_uint4(aIsLeaf ? (uint32_t)split : (uint32_t)split + n,
                                                         bIsLeaf ? (uint32_t)split + 1u : (uint32_t)split + 1u + n,
                                                         (uint32_t)range.x, cntA | (cntB << 16));
}

// ---- stage 4a: equal-key chains ---------------------------------------------------------------------------------------
// An equal-key run [s, e] of m >= 3 sorted positions is a right-leaning chain (BVH.h:153-154,212-233): chain node i
// (s < i < e) = (leaf i, node i+1), the last one (leaf e-1, leaf e).  Walking it bottom-up costs m dependent L2 round trips
// (ClumsyDragon: runs of up to 211, refit 0.40 ms of a 0.49 ms build).  The box of chain node i is the union of the leaf boxes
// i..e, i.e. a SUFFIX SCAN — one warp per run, shuffles, 32 positions per step.  It is bit-identical to the sequential walk:
// merge3 (AABB.h:107-113) is MIN(MIN(A, B), self) with the compare-select macros, whose ties (+0 / -0) go to the second
// operand, B = the suffix; "ties go to the right operand" is associative, so the scan keeps (lower positions, higher
// positions) operand order and gets the same bits.  The (FLT_MAX.., FLT_MIN..) start box clamps; clamping every step equals
// clamping the result.
struct Box6 { float lx, ly, lz, hx, hy, hz; };
__device__ __forceinline__ Box6 box_join(const Box6& L, const Box6& R) {      // L = lower sorted positions, R = higher
  Box6 o;
  o.lx = rmin(L.lx, R.lx); o.ly = rmin(L.ly, R.ly); o.lz = rmin(L.lz, R.lz);
  o.hx = rmax(L.hx, R.hx); o.hy = rmax(L.hy, R.hy); o.hz = rmax(L.hz, R.hz);
  return o;
}
__device__ __forceinline__ Box6 box_shfl_down(const Box6& b, int d) {
  Box6 o;
  o.lx = __shfl_down_sync(0xffffffffu, b.lx, d); o.ly = __shfl_down_sync(0xffffffffu, b.ly, d); o.lz = __shfl_down_sync(0xffffffffu, b.lz, d);
  o.hx = __shfl_down_sync(0xffffffffu, b.hx, d); o.hy = __shfl_down_sync(0xffffffffu, b.hy, d); o.hz = __shfl_down_sync(0xffffffffu, b.hz, d);
  return o;
}
__device__ __forceinline__ Box6 box_shfl(const Box6& b, int src) {
  Box6 o;
  o.lx = __shfl_sync(0xffffffffu, b.lx, src); o.ly = __shfl_sync(0xffffffffu, b.ly, src); o.lz = __shfl_sync(0xffffffffu, b.lz, src);
  o.hx = __shfl_sync(0xffffffffu, b.hx, src); o.hy = __shfl_sync(0xffffffffu, b.hy, src); o.hz = __shfl_sync(0xffffffffu, b.hz, src);
  return o;
}
__device__ __forceinline__ Box6 box_clamped(const Box6& b) {                  // the (FLT_MAX.., FLT_MIN..) start box takes part (Q4)
  Box6 o;
  o.lx = rmin(b.lx, RT_FLT_MAX); o.ly = rmin(b.ly, RT_FLT_MAX); o.lz = rmin(b.lz, RT_FLT_MAX);
  o.hx = rmax(b.hx, RT_FLT_MIN); o.hy = rmax(b.hy, RT_FLT_MIN); o.hz = rmax(b.hz, RT_FLT_MIN);
  return o;
}

__global__ void __launch_bounds__(256)
chain_suffix_kernel(const uint2* __restrict__ list, uint32_t n,
                    clrt_bvh_node* __restrict__ nodes, TraceNode* __restrict__ tnodes) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const uint32_t base = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 32u;
  if (base >= n) return;
  const uint32_t pos = base + lane;
  bool start = false;
  if (pos + 2 < n) {
    const uint32_t c = __ldg(&list[pos].x);
    start = __ldg(&list[pos + 1].x) == c && __ldg(&list[pos + 2].x) == c && (pos == 0 || __ldg(&list[pos - 1].x) != c);
  }
  unsigned starts = __ballot_sync(FULL, start);
  while (starts) {
    const uint32_t s = base + (uint32_t)(__ffs(starts) - 1);
    starts &= starts - 1;
    const uint32_t c = __ldg(&list[s].x);
    uint32_t e = s + 2;                                    // end of the run (inclusive), found 32 positions per step
    while (true) {
      const uint32_t q = e + 1 + lane;
      const unsigned same = __ballot_sync(FULL, q < n && __ldg(&list[q].x) == c);
      const int ext = same == FULL ? 32 : __ffs(~same) - 1;
      e += ext;
      if (ext < 32) break;
    }
    // chunks of 32 positions from the end of the run backwards; lane order = position order
    Box6 carry; carry.lx = carry.ly = carry.lz = carry.hx = carry.hy = carry.hz = 0.0f;
    bool have_carry = false;
    for (int64_t hi = e; hi >= (int64_t)s + 1; hi -= 32) {
      const int64_t p = hi - 31 + lane;
      const bool valid = p >= (int64_t)s + 1;
      Box6 v; v.lx = v.ly = v.lz = v.hx = v.hy = v.hz = 0.0f;
      if (valid) {
        const float4* lb = reinterpret_cast<const float4*>(nodes + __ldg(&list[p].y)) + 1;        // leaf box of morton_leaves_kernel
        const float4 lo = __ldcg(lb), hb = __ldcg(lb + 1);
        v.lx = lo.x; v.ly = lo.y; v.lz = lo.z; v.hx = hb.x; v.hy = hb.y; v.hz = hb.z;
      }
      const Box6 leaf = v;
#pragma unroll
      for (int d = 1; d < 32; d <<= 1) {
        const Box6 o = box_shfl_down(v, d);
        if (lane + d < 32) v = box_join(v, o);             // lanes below the run never feed a valid lane (suffix direction)
      }
      if (have_carry) v = box_join(v, carry);
      // child B of chain node p: leaf e itself for the last chain node, else chain node p+1
      Box6 next = box_shfl_down(v, 1);
      if (lane == 31) next = carry;
      const bool is_node = valid && p < (int64_t)e;        // position e is only a leaf
      if (is_node) {
        const Box6 mine = box_clamped(v);
        const Box6 b = (p + 1 == (int64_t)e) ? next : box_clamped(next);
        float4* po = reinterpret_cast<float4*>(nodes + n + p) + 1;
        __stcg(po, make_float4(mine.lx, mine.ly, mine.lz, 0.0f));
        __stcg(po + 1, make_float4(mine.hx, mine.hy, mine.hz, 0.0f));
        float4* tn = reinterpret_cast<float4*>(tnodes + p);
        __stcg(tn,     make_float4(leaf.lx, leaf.ly, leaf.lz, leaf.hx));
        __stcg(tn + 1, make_float4(leaf.hy, leaf.hz, b.lx, b.ly));
        __stcg(tn + 2, make_float4(b.lz, b.hx, b.hy, b.hz));
      }
      const int lowest = hi - 31 >= (int64_t)s + 1 ? 0 : (int)((int64_t)s + 1 - (hi - 31));
      carry = box_shfl(v, lowest);
      have_carry = true;
    }
  }
}

// ---- stage 4b: bottom-up refit with per-node arrival counters --------------------------------------------------------
// Leaves below a pre-computed chain (positions s+2..e of a run of 3+) have nothing to do; the leaf at s+1 stands in for the
// whole chain and arrives at the chain's parent as its child B.
__global__ void __launch_bounds__(256)
refit_atomic_kernel(clrt_bvh_node* __restrict__ nodes, uint32_t* __restrict__ counters, uint32_t n, TraceNode* __restrict__ tnodes,
                    const uint2* __restrict__ list, const uint32_t* __restrict__ refpos) {
  const uint32_t gid = blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= n) return;
  // Thread = SORTED POSITION, not triangle id: the lanes of a warp are neighbours in the tree, so the arrival counters, parent
  // headers and boxes they touch share cache lines (round 1 mapped threads to triangle ids, i.e. every lane walked an unrelated path:
  // 315 MB of DRAM traffic for 100 MB algorithmic).
  uint32_t current, me;
  {
    const uint32_t pos = gid;
    const uint2 key = __ldg(list + pos);
    me = key.y;
    const uint32_t c = key.x;
    const bool prevEq = pos > 0 && __ldg(&list[pos - 1].x) == c;
    const bool prev2Eq = prevEq && pos > 1 && __ldg(&list[pos - 2].x) == c;
    const bool nextEq = pos + 1 < n && __ldg(&list[pos + 1].x) == c;
    if (prev2Eq) return;                                                     // s+2..e: inside a pre-computed chain
    if (prevEq && nextEq) me = n + pos;                                      // s+1 of a run of 3+: stands in for chain node s+1
    current = __ldcg(&nodes[me].data[0]);
  }
  // The thread carries the box of the subtree it comes from (`me`) in registers; per level it needs ONE dependent load (the
  // sibling's box) after the arrival atomic — the parent's header (children, grandparent) is static and is fetched while the
  // fence and the atomic are in flight.
  float4 mlo = ldcg4(reinterpret_cast<const float4*>(nodes + me) + 1), mhi = ldcg4(reinterpret_cast<const float4*>(nodes + me) + 2);
  while (current != 0xFFFFFFFFu) {
    const uint4 d = __ldcg(reinterpret_cast<const uint4*>(nodes + current));   // {parent, childA, childB, -}: written by the emit stage
    // release: this thread's box (written below / by the leaf stage) is visible before the count; acquire: the sibling's box,
    // written before ITS increment, is visible after ours (Q3: the reference has neither and races)
    uint32_t arrived;
    asm volatile("atom.add.acq_rel.gpu.global.u32 %0, [%1], 1;" : "=r"(arrived) : "l"(counters + (current - n)) : "memory");
    if (arrived == 0) return;                                // first arrival: the sibling subtree is not done yet
    const bool fromA = d.y == me;
    const float4* ps = reinterpret_cast<const float4*>(nodes + (fromA ? d.z : d.y)) + 1;
    const float4 s0 = ldcg4(ps), s1 = ldcg4(ps + 1);
    const float4 a0 = fromA ? mlo : s0, a1 = fromA ? mhi : s1, b0 = fromA ? s0 : mlo, b1 = fromA ? s1 : mhi;
    // merge3(childA, childB, self) (AABB.h:107-113) with self = (FLT_MAX,FLT_MAX,FLT_MAX,INNER=0.0f),(FLT_MIN,FLT_MIN,FLT_MIN,0.0f)
    float4 lo, hi;
    lo.x = rmin(rmin(a0.x, b0.x), RT_FLT_MAX); lo.y = rmin(rmin(a0.y, b0.y), RT_FLT_MAX); lo.z = rmin(rmin(a0.z, b0.z), RT_FLT_MAX);
    lo.w = rmin(rmin(a0.w, b0.w), 0.0f);
    hi.x = rmax(rmax(a1.x, b1.x), RT_FLT_MIN); hi.y = rmax(rmax(a1.y, b1.y), RT_FLT_MIN); hi.z = rmax(rmax(a1.z, b1.z), RT_FLT_MIN);
    hi.w = rmax(rmax(a1.w, b1.w), 0.0f);
    float4* po = reinterpret_cast<float4*>(nodes + current) + 1;
    __stcg(po, lo); __stcg(po + 1, hi);
    float4* tn = reinterpret_cast<float4*>(tnodes + (current - n));
    __stcg(tn,     make_float4(a0.x, a0.y, a0.z, a1.x));
    __stcg(tn + 1, make_float4(a1.y, a1.z, b0.x, b0.y));
    __stcg(tn + 2, make_float4(b0.z, b1.x, b1.y, b1.z));
    mlo = lo; mhi = hi; me = current;
    current = d.x;
  }
}

} // namespace

// ---- manager ---------------------------------------------------------------------------------------------------
int bvh_init_frame(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  const uint32_t n = a->scene->n_tris;
  a->n = n;
  a->built = false;
  a->tt_built = false;
  a->pairs_items = pow2ceil_u32(n ? n : 1);                  // the reference pads to 2^k for its bitonic sort (BVHManager.cpp:129-131)
  const size_t nodes_bytes = (size_t)(n ? 2 * (size_t)n - 1 : 1) * sizeof(clrt_bvh_node);
  if (nodes_bytes > a->nodes_cap || (size_t)a->pairs_items * 8 > a->pairs_cap || (size_t)n * 4 > a->counters_cap ||
      (size_t)n * 4 > a->refpos_cap || (size_t)n * sizeof(TraceNode) > a->tnodes_cap || (size_t)n * sizeof(clrt_triangle) > a->tris_sorted_cap)
    CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
  CLRT_TRY(clrt_grow((void**)&a->d_nodes, &a->nodes_cap, nodes_bytes));
  CLRT_TRY(clrt_grow((void**)&a->d_pairs, &a->pairs_cap, (size_t)a->pairs_items * 8));
  CLRT_TRY(clrt_grow((void**)&a->d_counters, &a->counters_cap, (size_t)(n ? n : 1) * 4));
  CLRT_TRY(clrt_grow((void**)&a->d_refpos, &a->refpos_cap, (size_t)(n ? n : 1) * 4));
  CLRT_TRY(clrt_grow((void**)&a->d_tnodes, &a->tnodes_cap, (size_t)(n ? n : 1) * sizeof(TraceNode)));
  CLRT_TRY(clrt_grow((void**)&a->d_tris_sorted, &a->tris_sorted_cap, (size_t)(n ? n : 1) * sizeof(clrt_triangle)));
  // fills of initializeFrame (BVHManager.cpp:138-147): pairs = (UINT_MAX,UINT_MAX), counters = 0
  CLRT_CUDA(cudaMemsetAsync(a->d_pairs, 0xFF, (size_t)a->pairs_items * 8, ctx->stream));
  CLRT_CUDA(cudaMemsetAsync(a->d_counters, 0, (size_t)(n ? n : 1) * 4, ctx->stream));
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
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter + 12, 0, 4, s));
  if (n) {
    karras_emit_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->d_pairs, n, a->d_nodes, a->scene->d_tris, a->d_tris_sorted, a->d_tnodes, a->d_tile_counter + 12, a->d_refpos);
    CLRT_CHECK_LAUNCH();
  }
  if (tm) CLRT_CUDA(cudaEventRecord(a->ev[3], s));
  if (n > 1) {
    chain_suffix_kernel<<<(n + 255) / 256, 256, 0, s>>>(reinterpret_cast<const uint2*>(a->d_pairs), n, a->d_nodes, a->d_tnodes);
    CLRT_CHECK_LAUNCH();
    refit_atomic_kernel<<<(n + 255) / 256, 256, 0, s>>>(a->d_nodes, a->d_counters, n, a->d_tnodes, reinterpret_cast<const uint2*>(a->d_pairs), a->d_refpos);
    CLRT_CHECK_LAUNCH();
  }
  if (tm) {
    CLRT_CUDA(cudaEventRecord(a->ev[4], s));
    CLRT_CUDA(cudaEventSynchronize(a->ev[4]));
    for (int i = 0; i < 4; i++) CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[i], a->ev[i], a->ev[i + 1]));
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
