// tlg.cu — Two-Level Grid (Kalojanov et al.) builder + 3D-DDA traversal.
// Replaces TwoLevelGridManager::initializeFrame / construct / generateContacts
// (R/src/CLRayTracerEngine/TwoLevelGridManager.cpp:160-201, 207-469, 475-560), the kernels of
// R/src/CLRayTracerEngine/TwoLevelGridKernels.cl and the device functions of
// R/Include/CLData/AccelerationStructs/TwoLevelGrid.h.
//
// Build sequence (same stages as the reference; no padding to powers of two, no per-stage host wait):
//   tlg_count_top     <- prepareDataKernel (.cl:43-51 -> TwoLevelGrid.h:87-124)        48 B tri in, 4 B out
//   scan              <- PrefixSum                                                       8 B / element
//   tlg_scatter_top   <- writePairsKernel (.cl:56-66 -> :136-190)                        8 B / pair
//   sort              <- BitonicSort                                                    16 B / pair
//   extract_ranges    <- extractCellRangesKernel (.cl:71-116)                            8 B / pair in, 8 B / boundary out
//   tlg_leaf_res      <- countLeavesAndFillCellKernel (.cl:121-127 -> :199-225)          8 B in, 20 B out / cell
//   scan, tlg_first_leaf <- updateTopLevelCellsWithLeafRange (.cl:132-139)
//   tlg_count_leaf    <- prepareGridDataForLeaves (.cl:144-157 -> :252-315)              8 B pair + 48 B tri + 16 B cell in, 4 B out
//   scan, tlg_scatter_leaf <- writeLeafPairsKernel (.cl:162-174 -> :330-403, AABB.h:460-540 exact SAT)
//   sort, extract_ranges   <- extractLeafCellsKernel (.cl:179-221)
// Three small read-backs remain (pair counts size the next allocation), as in the reference (:237,:345,:393).
//
// Parity stance: Q17 FIXED (every triangle takes part), Q18 FIXED (range[c] = [first, one past
// last) over the real pairs; empty = (0,0)), Q19 LITERAL (conservative leaf counts leave gaps that sort last), Q20 (pow in
// double, rounded once), Q21 LITERAL traversal arithmetic over zero-initialised buffers, Q22 OpenCL sign().
// All arithmetic keeps the reference's operation order (-fmad=false), so cell lists and contacts are bit-exact.
//
// The per-pair loops of the reference run over EVERY leaf cell of a top cell (TwoLevelGrid.h:269-280,351-364); here they
// run over the leaf-index window the triangle's box can reach (widened by one cell + a rounding allowance), inside
// which the reference's exact tests decide — same pairs, same order (z, y, x ascending).
#include "common.cuh"
#include "rtmath.cuh"
#include <math.h>

struct TlgState {
  clrt_grid_data grid;
  uint32_t cells = 0, top_pairs = 0, leaf_cells = 0, leaf_pairs_cap = 0, leaf_pairs = 0;
  uint32_t* d_counters = nullptr; size_t counters_cap = 0;
  uint32_t* d_scan = nullptr; size_t scan_cap = 0;
  clrt_uint2* d_top_ranges = nullptr; size_t top_ranges_cap = 0;
  clrt_top_cell* d_top_cells = nullptr; size_t top_cells_cap = 0;
  clrt_uint2* d_top_pairs = nullptr; size_t top_pairs_cap = 0;
  clrt_uint2* d_leaf_pairs = nullptr; size_t leaf_pairs_capacity = 0;
  clrt_uint2* d_leaf_ranges = nullptr; size_t leaf_ranges_cap = 0;
  uint32_t* d_real = nullptr;     // [0] = number of real leaf pairs (set by extract_ranges)
  bool built = false;
};

namespace {

constexpr float ONE_THIRD = 1.0f / 3.0f;          // TwoLevelGrid.h:192
constexpr uint32_t GAP = 0xFFFFFFFFu;

__device__ __forceinline__ void load_tri(const clrt_triangle* __restrict__ tris, uint32_t i, Vec3& a, Vec3& b, Vec3& c) {
  const float4* t = reinterpret_cast<const float4*>(tris + i);
  const float4 q0 = __ldg(t), q1 = __ldg(t + 1), q2 = __ldg(t + 2);
  a = v3(q0.x, q0.y, q0.z); b = v3(q1.x, q1.y, q1.z); c = v3(q2.x, q2.y, q2.z);
}

// getCellRefFromIndex (TwoLevelGrid.h:69-75)
__device__ __forceinline__ void cell_ref(uint32_t idx, uint32_t rx, uint32_t ry, uint32_t& x, uint32_t& y, uint32_t& z) {
  const uint32_t a = rx * ry; z = idx / a; const uint32_t b = idx - a * z; x = b % rx; y = b / rx;
}

// ---- top level -----------------------------------------------------------------------------------------------------
// countOverlappingCells (TwoLevelGrid.h:87-100): float arithmetic throughout, product converted to uint.
__device__ __forceinline__ uint32_t count_top(Vec3 v0, Vec3 v1, Vec3 v2, const clrt_grid_data& g) {
  const float lx = (float)(g.resX - 1), ly = (float)(g.resY - 1), lz = (float)(g.resZ - 1);
  const float mnx = rmin(v0.x, rmin(v1.x, v2.x)), mny = rmin(v0.y, rmin(v1.y, v2.y)), mnz = rmin(v0.z, rmin(v1.z, v2.z));
  const float mxx = rmax(v0.x, rmax(v1.x, v2.x)), mxy = rmax(v0.y, rmax(v1.y, v2.y)), mxz = rmax(v0.z, rmax(v1.z, v2.z));
  const float sx = rmin(floorf((mnx - g.box.bounds[0].x) / g.stepX), lx);
  const float sy = rmin(floorf((mny - g.box.bounds[0].y) / g.stepY), ly);
  const float sz = rmin(floorf((mnz - g.box.bounds[0].z) / g.stepZ), lz);
  const float ex = rmin(floorf((mxx - g.box.bounds[0].x) / g.stepX), lx);
  const float ey = rmin(floorf((mxy - g.box.bounds[0].y) / g.stepY), ly);
  const float ez = rmin(floorf((mxz - g.box.bounds[0].z) / g.stepZ), lz);
  const float cx = (ex - sx) + 1.0f, cy = (ey - sy) + 1.0f, cz = (ez - sz) + 1.0f;
  return (uint32_t)(cx * cy * cz);
}

__global__ void __launch_bounds__(256)
tlg_count_top_kernel(const clrt_triangle* __restrict__ tris, uint32_t n, const clrt_grid_data g, uint32_t* __restrict__ counters) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Vec3 a, b, c; load_tri(tris, i, a, b, c);
  counters[i] = count_top(a, b, c, g);
}

// writeOverlappingPairs (TwoLevelGrid.h:136-163): uint cell window, pairs at scan[i]-count[i] in z,y,x order.
__global__ void __launch_bounds__(256)
tlg_scatter_top_kernel(const clrt_triangle* __restrict__ tris, uint32_t n, const clrt_grid_data g,
                       const uint32_t* __restrict__ scan, const uint32_t* __restrict__ counters, uint32_t cap, uint2* __restrict__ pairs) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Vec3 v0, v1, v2; load_tri(tris, i, v0, v1, v2);
  const float mnx = rmin(v0.x, rmin(v1.x, v2.x)), mny = rmin(v0.y, rmin(v1.y, v2.y)), mnz = rmin(v0.z, rmin(v1.z, v2.z));
  const float mxx = rmax(v0.x, rmax(v1.x, v2.x)), mxy = rmax(v0.y, rmax(v1.y, v2.y)), mxz = rmax(v0.z, rmax(v1.z, v2.z));
  const uint32_t mx = g.resX - 1, my = g.resY - 1, mz = g.resZ - 1;
  const uint32_t x0 = min((uint32_t)floorf((mnx - g.box.bounds[0].x) / g.stepX), mx), x1 = min((uint32_t)floorf((mxx - g.box.bounds[0].x) / g.stepX), mx);
  const uint32_t y0 = min((uint32_t)floorf((mny - g.box.bounds[0].y) / g.stepY), my), y1 = min((uint32_t)floorf((mxy - g.box.bounds[0].y) / g.stepY), my);
  const uint32_t z0 = min((uint32_t)floorf((mnz - g.box.bounds[0].z) / g.stepZ), mz), z1 = min((uint32_t)floorf((mxz - g.box.bounds[0].z) / g.stepZ), mz);
  uint32_t at = scan[i] - counters[i];
  for (uint32_t z = z0; z <= z1; z++)
    for (uint32_t y = y0; y <= y1; y++)
      for (uint32_t x = x0; x <= x1; x++, at++)
        if (at < cap) pairs[at] = make_uint2(z * g.resX * g.resY + y * g.resX + x, i);
}

// extractCellRangesKernel / extractLeafCellsKernel with the FIXED semantics (Q18): over the real pairs (key != GAP)
// range[c] = [first index, one past last index); also reports the number of real pairs.
__global__ void __launch_bounds__(256)
extract_ranges_kernel(const uint2* __restrict__ sorted, uint32_t count, uint2* __restrict__ ranges, uint32_t* __restrict__ real_count) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= count) return;
  const uint32_t me = sorted[i].x;
  if (me == GAP) return;
  const uint32_t right = (i + 1 < count) ? sorted[i + 1].x : GAP;
  if (i == 0) ranges[me].x = 0;
  if (right > me) {
    ranges[me].y = i + 1;
    if (right != GAP) ranges[right].x = i + 1;
    else if (real_count) *real_count = i + 1;
  }
}

// ---- leaf level ----------------------------------------------------------------------------------------------------
// calcLeafCellResolution (TwoLevelGrid.h:199-205) + fillTopLevelCell (:215-225).  pow: double, rounded once (Q20).
__global__ void __launch_bounds__(256)
tlg_leaf_res_kernel(const uint2* __restrict__ ranges, uint32_t cells, const clrt_grid_data g,
                    clrt_top_cell* __restrict__ top, uint32_t* __restrict__ leaf_counts) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells) return;
  const uint2 r = ranges[c];
  const uint32_t prims = r.y - r.x;
  const float volume = g.stepX * g.stepY * g.stepZ;
  const float a = (float)pow((double)(g.leafDensity * (float)prims / volume), (double)ONE_THIRD);
  const uint32_t rx = (uint32_t)floorf(g.stepX * a), ry = (uint32_t)floorf(g.stepY * a), rz = (uint32_t)floorf(g.stepZ * a);
  reinterpret_cast<uint4*>(top)[c] = make_uint4(rx, ry, rz, 0u);
  leaf_counts[c] = rx * ry * rz;
}
// updateTopLevelCellsWithLeafRange (.cl:132-139): firstLeafIdx[c] = inclusiveScan[c-1]
__global__ void __launch_bounds__(256)
tlg_first_leaf_kernel(clrt_top_cell* __restrict__ top, const uint32_t* __restrict__ scan, uint32_t cells) {
  const uint32_t c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cells || c == 0) return;
  top[c].firstLeafIdx = scan[c - 1];
}

struct LeafFrame { float stepX, stepY, stepZ, baseX, baseY, baseZ; uint32_t x0, x1, y0, y1, z0, z1; };

// Leaf-cell frame of a top cell (TwoLevelGrid.h:259-266 / :337-344) and the index window the triangle's box can reach.
__device__ __forceinline__ LeafFrame leaf_frame(const clrt_grid_data& g, const uint4 cell, uint32_t cellIdx,
                                                float mnx, float mny, float mnz, float mxx, float mxy, float mxz) {
  LeafFrame f;
  f.stepX = g.stepX / (float)cell.x; f.stepY = g.stepY / (float)cell.y; f.stepZ = g.stepZ / (float)cell.z;
  uint32_t cx, cy, cz; cell_ref(cellIdx, g.resX, g.resY, cx, cy, cz);
  f.baseX = g.box.bounds[0].x + (float)cx * g.stepX;
  f.baseY = g.box.bounds[0].y + (float)cy * g.stepY;
  f.baseZ = g.box.bounds[0].z + (float)cz * g.stepZ;
  // window: floor((tri - base)/step) widened by one cell and by a rounding allowance proportional to the coordinates
  auto lo = [](float v, float base, float step, uint32_t res) -> uint32_t {
    const float slack = 8.0f * RT_FLT_EPSILON * (fabsf(v) + fabsf(base) + step * (float)res);
    const float k = floorf((v - slack - base) / step) - 1.0f;
    return k > 0.0f ? (k < (float)res ? (uint32_t)k : res) : 0u;
  };
  auto hi = [](float v, float base, float step, uint32_t res) -> uint32_t {       // inclusive upper index, or res-1
    const float slack = 8.0f * RT_FLT_EPSILON * (fabsf(v) + fabsf(base) + step * (float)res);
    const float k = floorf((v + slack - base) / step) + 1.0f;
    return k < 0.0f ? 0u : (k < (float)(res - 1) ? (uint32_t)k : res - 1);
  };
  f.x0 = lo(mnx, f.baseX, f.stepX, cell.x); f.x1 = hi(mxx, f.baseX, f.stepX, cell.x);
  f.y0 = lo(mny, f.baseY, f.stepY, cell.y); f.y1 = hi(mxy, f.baseY, f.stepY, cell.y);
  f.z0 = lo(mnz, f.baseZ, f.stepZ, cell.z); f.z1 = hi(mxz, f.baseZ, f.stepZ, cell.z);
  return f;
}

// countOverlappingLeafCells (TwoLevelGrid.h:252-282): AABBOverlaps (AABB.h:205-212) of each leaf cell with
// calculateTriangleAABB (AABB.h:61-79, flat boxes padded by FLT_EPSILON).
__global__ void __launch_bounds__(128)
tlg_count_leaf_kernel(const clrt_triangle* __restrict__ tris, const uint2* __restrict__ top_pairs, uint32_t pairs,
                      const clrt_grid_data g, const clrt_top_cell* __restrict__ top, uint32_t* __restrict__ counters) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= pairs) return;
  const uint2 pr = top_pairs[p];
  Vec3 v1, v2, v3_; load_tri(tris, pr.y, v1, v2, v3_);
  float mnx = rmin(rmin(v1.x, v2.x), v3_.x), mny = rmin(rmin(v1.y, v2.y), v3_.y), mnz = rmin(rmin(v1.z, v2.z), v3_.z);
  float mxx = rmax(rmax(v1.x, v2.x), v3_.x), mxy = rmax(rmax(v1.y, v2.y), v3_.y), mxz = rmax(rmax(v1.z, v2.z), v3_.z);
  const float xF = (float)(uint32_t)((mxx - mnx) < RT_FLT_EPSILON), yF = (float)(uint32_t)((mxy - mny) < RT_FLT_EPSILON), zF = (float)(uint32_t)((mxz - mnz) < RT_FLT_EPSILON);
  mxx += xF * RT_FLT_EPSILON; mxy += yF * RT_FLT_EPSILON; mxz += zF * RT_FLT_EPSILON;
  mnx -= xF * RT_FLT_EPSILON; mny -= yF * RT_FLT_EPSILON; mnz -= zF * RT_FLT_EPSILON;
  const uint4 cell = __ldg(reinterpret_cast<const uint4*>(top) + pr.x);
  uint32_t count = 0;
  if (cell.x && cell.y && cell.z) {
    const LeafFrame f = leaf_frame(g, cell, pr.x, mnx, mny, mnz, mxx, mxy, mxz);
    for (uint32_t z = f.z0; z <= f.z1; z++) {
      const float bz0 = f.baseZ + (float)z * f.stepZ, bz1 = bz0 + f.stepZ;
      const bool zNo = rmin(bz1, mxz) < rmax(bz0, mnz);
      for (uint32_t y = f.y0; y <= f.y1; y++) {
        const float by0 = f.baseY + (float)y * f.stepY, by1 = by0 + f.stepY;
        const bool yNo = rmin(by1, mxy) < rmax(by0, mny);
        for (uint32_t x = f.x0; x <= f.x1; x++) {
          const float bx0 = f.baseX + (float)x * f.stepX, bx1 = bx0 + f.stepX;
          const bool xNo = rmin(bx1, mxx) < rmax(bx0, mnx);
          count += (uint32_t)!(xNo || yNo || zNo);
        }
      }
    }
  }
  counters[p] = count;
}

// OpenCL sign() (Q22: the host macro of CLPortability.h:282 has a precedence bug; the device runs the builtin)
__device__ __forceinline__ float cl_sign(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : (v == 0.0f ? v : 0.0f)); }

// AABBTriangleIntersect (AABB.h:460-540; Akenine-Moller tribox3) with planeBoxOverlap (:436-455), literal order.
__device__ __forceinline__ bool aabb_triangle_intersect(Vec3 boxcenter, Vec3 h, Vec3 v0, Vec3 v1, Vec3 v2) {
  float minimum, maximum, p0, p1, p2, rad;
  bool ok = true;
  v0 = vsub(boxcenter, v0); v1 = vsub(boxcenter, v1); v2 = vsub(boxcenter, v2);
  {
    Vec3 e = vsub(v1, v0);
    float fex = fabsf(e.x), fey = fabsf(e.y), fez = fabsf(e.z);
    // AXISTEST_X01(e.z, e.y, fez, fey)
    p0 = e.z * v0.y - e.y * v0.z; p2 = e.z * v2.y - e.y * v2.z; minimum = rmin(p0, p2); maximum = rmax(p0, p2);
    rad = fez * h.y + fey * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    // AXISTEST_Y02(e.z, e.x, fez, fex)
    p0 = -e.z * v0.x + e.x * v0.z; p2 = -e.z * v2.x + e.x * v2.z; minimum = rmin(p0, p2); maximum = rmax(p0, p2);
    rad = fez * h.x + fex * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    // AXISTEST_Z12(e.y, e.x, fey, fex)
    p1 = e.y * v1.x - e.x * v1.y; p2 = e.y * v2.x - e.x * v2.y; minimum = rmin(p1, p2); maximum = rmax(p1, p2);
    rad = fey * h.x + fex * h.y; ok = ok && (minimum <= rad && maximum >= -rad);

    e = vsub(v2, v1); fex = fabsf(e.x); fey = fabsf(e.y); fez = fabsf(e.z);
    p0 = e.z * v0.y - e.y * v0.z; p2 = e.z * v2.y - e.y * v2.z; minimum = rmin(p0, p2); maximum = rmax(p0, p2);
    rad = fez * h.y + fey * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    p0 = -e.z * v0.x + e.x * v0.z; p2 = -e.z * v2.x + e.x * v2.z; minimum = rmin(p0, p2); maximum = rmax(p0, p2);
    rad = fez * h.x + fex * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    // AXISTEST_Z0(e.y, e.x, fey, fex)
    p0 = e.y * v0.x - e.x * v0.y; p1 = e.y * v1.x - e.x * v1.y; minimum = rmin(p0, p1); maximum = rmax(p0, p1);
    rad = fey * h.x + fex * h.y; ok = ok && (minimum <= rad && maximum >= -rad);

    e = vsub(v0, v2); fex = fabsf(e.x); fey = fabsf(e.y); fez = fabsf(e.z);
    // AXISTEST_X2(e.z, e.y, fez, fey)
    p0 = e.z * v0.y - e.y * v0.z; p1 = e.z * v1.y - e.y * v1.z; minimum = rmin(p0, p1); maximum = rmax(p0, p1);
    rad = fez * h.y + fey * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    // AXISTEST_Y1(e.z, e.x, fez, fex)
    p0 = -e.z * v0.x + e.x * v0.z; p1 = -e.z * v1.x + e.x * v1.z; minimum = rmin(p0, p1); maximum = rmax(p0, p1);
    rad = fez * h.x + fex * h.z; ok = ok && (minimum <= rad && maximum >= -rad);
    // AXISTEST_Z12(e.y, e.x, fey, fex)
    p1 = e.y * v1.x - e.x * v1.y; p2 = e.y * v2.x - e.x * v2.y; minimum = rmin(p1, p2); maximum = rmax(p1, p2);
    rad = fey * h.x + fex * h.y; ok = ok && (minimum <= rad && maximum >= -rad);
  }
  minimum = rmin(v0.x, rmin(v1.x, v2.x)); maximum = rmax(v0.x, rmax(v1.x, v2.x)); ok = ok && (minimum <= h.x && maximum >= -h.x);
  minimum = rmin(v0.y, rmin(v1.y, v2.y)); maximum = rmax(v0.y, rmax(v1.y, v2.y)); ok = ok && (minimum <= h.y && maximum >= -h.y);
  minimum = rmin(v0.z, rmin(v1.z, v2.z)); maximum = rmax(v0.z, rmax(v1.z, v2.z)); ok = ok && (minimum <= h.z && maximum >= -h.z);
  const Vec3 normal = vcross(vsub(v1, v0), vsub(v2, v1));
  const float sx = cl_sign(normal.x), sy = cl_sign(normal.y), sz = cl_sign(normal.z);
  const Vec3 vmin = v3(-sx * h.x - v0.x, -sy * h.y - v0.y, -sz * h.z - v0.z);
  const Vec3 vmax = v3(sx * h.x - v0.x, sy * h.y - v0.y, sz * h.z - v0.z);
  bool plane = vdot(normal, vmin) <= 0.0f;
  plane = plane && vdot(normal, vmax) >= 0.0f;
  return ok && plane;
}

// writeOverlappingLeafPairs (TwoLevelGrid.h:330-365): exact test, pairs at scan[p]-count[p] in z,y,x order.
__global__ void __launch_bounds__(128)
tlg_scatter_leaf_kernel(const clrt_triangle* __restrict__ tris, const uint2* __restrict__ top_pairs, uint32_t pairs,
                        const clrt_grid_data g, const clrt_top_cell* __restrict__ top, const uint32_t* __restrict__ scan,
                        const uint32_t* __restrict__ counters, uint32_t cap, uint2* __restrict__ leaf_pairs, uint32_t* __restrict__ over_budget) {
  const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= pairs) return;
  const uint32_t budget = counters[p];
  if (!budget) return;
  const uint2 pr = top_pairs[p];
  Vec3 v0, v1, v2; load_tri(tris, pr.y, v0, v1, v2);
  const float mnx = rmin(v0.x, rmin(v1.x, v2.x)), mny = rmin(v0.y, rmin(v1.y, v2.y)), mnz = rmin(v0.z, rmin(v1.z, v2.z));
  const float mxx = rmax(v0.x, rmax(v1.x, v2.x)), mxy = rmax(v0.y, rmax(v1.y, v2.y)), mxz = rmax(v0.z, rmax(v1.z, v2.z));
  const uint4 cell = __ldg(reinterpret_cast<const uint4*>(top) + pr.x);
  const LeafFrame f = leaf_frame(g, cell, pr.x, mnx, mny, mnz, mxx, mxy, mxz);
  const Vec3 half = v3(f.stepX * 0.5f, f.stepY * 0.5f, f.stepZ * 0.5f);
  uint32_t at = scan[p] - budget, written = 0;
  for (uint32_t z = f.z0; z <= f.z1; z++)
    for (uint32_t y = f.y0; y <= f.y1; y++)
      for (uint32_t x = f.x0; x <= f.x1; x++) {
        const Vec3 centre = v3(f.baseX + (float)x * f.stepX + half.x, f.baseY + (float)y * f.stepY + half.y, f.baseZ + (float)z * f.stepZ + half.z);
        if (aabb_triangle_intersect(centre, half, v0, v1, v2)) {
          // The budget comes from the conservative box count (tlg_count_leaf_kernel), the writes from the exact test: should rounding
          // ever let the exact test accept a cell the count did not see, the pair must not land in the next top pair's region (the
          // reference has the same hole, TwoLevelGrid.h:330-365, and races there).  It is dropped and COUNTED: construct() fails.
          if (written < budget && at + written < cap) leaf_pairs[at + written] = make_uint2(z * cell.x * cell.y + y * cell.x + x + cell.w, pr.y);
          else atomicAdd(over_budget, 1u);
          written++;
        }
      }
}

// ---- traversal -------------------------------------------------------------------------------------------------------
// tlg_generate_contact (TwoLevelGrid.h:557-678) + processTopLevelCell (:419-545), literal arithmetic (Q21): entry
// t0 = min of the three slab entries, first leaf cell with any hit ends the walk, 1.0/dir.z evaluated in double.
// Differences: triangles come from the flattened array; a step cap guarantees termination where the reference's
// `while (minimal != next[axis]) axis++` would run away on NaN.
struct GridHit { float t; uint32_t tri; };

struct Dda {
  float nx, ny, nz, dx, dy, dz; int ix, iy, iz, sx, sy, sz, ex, ey, ez;
};

__device__ __forceinline__ float clampf_macro(float v, float l, float h) { return rmin(rmax(v, l), h); }   // CLPortability.h:401

// Common set-up of both levels: box [b0,b1], resolution r.  Returns false when the top-level early-out fires (top only).
__device__ __forceinline__ bool dda_setup(Dda& s, Vec3 o, Vec3 d, Vec3 inv, float b0x, float b0y, float b0z, float b1x, float b1y, float b1z,
                                          uint32_t rx, uint32_t ry, uint32_t rz, bool top_level) {
  const float tx_min = ((inv.x >= 0.0f ? b0x : b1x) - o.x) * inv.x, tx_max = ((inv.x >= 0.0f ? b1x : b0x) - o.x) * inv.x;
  const float ty_min = ((inv.y >= 0.0f ? b0y : b1y) - o.y) * inv.y, ty_max = ((inv.y >= 0.0f ? b1y : b0y) - o.y) * inv.y;
  const float tz_min = ((inv.z >= 0.0f ? b0z : b1z) - o.z) * inv.z, tz_max = ((inv.z >= 0.0f ? b1z : b0z) - o.z) * inv.z;
  const float t0 = rmin(tx_min, rmin(ty_min, tz_min));
  if (top_level && t0 > rmax(tx_max, rmax(ty_max, tz_max))) return false;
  const bool inside = point_inside(b0x, b0y, b0z, b1x, b1y, b1z, o);
  const float px = inside ? o.x : o.x + d.x * t0, py = inside ? o.y : o.y + d.y * t0, pz = inside ? o.z : o.z + d.z * t0;
  s.ix = (int)clampf_macro((px - b0x) * (float)rx / (b1x - b0x), 0.0f, (float)rx - 1);
  s.iy = (int)clampf_macro((py - b0y) * (float)ry / (b1y - b0y), 0.0f, (float)ry - 1);
  s.iz = (int)clampf_macro((pz - b0z) * (float)rz / (b1z - b0z), 0.0f, (float)rz - 1);
  s.dx = (tx_max - tx_min) / (float)rx; s.dy = (ty_max - ty_min) / (float)ry; s.dz = (tz_max - tz_min) / (float)rz;
  if (d.x > 0.0f) { s.nx = tx_min + (float)(s.ix + 1) * s.dx; s.sx = 1; s.ex = (int)rx; }
  else { s.nx = d.x == 0.0f ? RT_FLT_MAX : tx_min + (float)(rx - (uint32_t)s.ix) * s.dx; s.sx = -1; s.ex = -1; }
  if (d.y > 0.0f) { s.ny = ty_min + (float)(s.iy + 1) * s.dy; s.sy = 1; s.ey = (int)ry; }
  else { s.ny = d.y == 0.0f ? RT_FLT_MAX : ty_min + (float)(ry - (uint32_t)s.iy) * s.dy; s.sy = -1; s.ey = -1; }
  if (d.z > 0.0f) { s.nz = tz_min + (float)(s.iz + 1) * s.dz; s.sz = 1; s.ez = (int)rz; }
  else { s.nz = d.z == 0.0f ? RT_FLT_MAX : tz_min + (float)(rz - (uint32_t)s.iz) * s.dz; s.sz = -1; s.ez = -1; }
  return true;
}
// one DDA step; returns false when the walk leaves the grid
__device__ __forceinline__ bool dda_step(Dda& s, int axis) {
  if (axis == 0) { s.nx += s.dx; s.ix += s.sx; return s.ix != s.ex; }
  if (axis == 1) { s.ny += s.dy; s.iy += s.sy; return s.iy != s.ey; }
  s.nz += s.dz; s.iz += s.sz; return s.iz != s.ez;
}
__device__ __forceinline__ int dda_axis(const Dda& s) {
  const float minimal = rmin(s.nx, rmin(s.ny, s.nz));
  return minimal == s.nx ? 0 : (minimal == s.ny ? 1 : 2);
}

__device__ __forceinline__ GridHit tlg_traverse(Vec3 o, Vec3 d, const clrt_grid_data& g, const clrt_top_cell* __restrict__ top,
                                                const uint2* __restrict__ leaf_ranges, const uint2* __restrict__ leaf_pairs,
                                                const clrt_triangle* __restrict__ tris) {
  GridHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
  const Vec3 inv = v3(1.0f / d.x, 1.0f / d.y, (float)(1.0 / (double)d.z));
  Dda T;
  if (!dda_setup(T, o, d, inv, g.box.bounds[0].x, g.box.bounds[0].y, g.box.bounds[0].z, g.box.bounds[1].x, g.box.bounds[1].y, g.box.bounds[1].z,
                 g.resX, g.resY, g.resZ, true)) return none;
  const uint32_t top_cap = g.resX + g.resY + g.resZ + 3u;
  for (uint32_t it = 0; it < top_cap; it++) {
    const uint4 cell = __ldg(reinterpret_cast<const uint4*>(top) + ((uint32_t)T.iz * g.resX * g.resY + (uint32_t)T.iy * g.resX + (uint32_t)T.ix));
    const int axis = dda_axis(T);
    if (!(cell.x == 0 || cell.y == 0 || cell.z == 0)) {
      const float b0x = g.box.bounds[0].x + (float)T.ix * g.stepX, b0y = g.box.bounds[0].y + (float)T.iy * g.stepY, b0z = g.box.bounds[0].z + (float)T.iz * g.stepZ;
      Dda Lf;
      dda_setup(Lf, o, d, inv, b0x, b0y, b0z, b0x + g.stepX, b0y + g.stepY, b0z + g.stepZ, cell.x, cell.y, cell.z, false);
      GridHit h; h.t = RT_FLT_MAX; h.tri = 0xFFFFFFFFu;
      const uint32_t leaf_cap = cell.x + cell.y + cell.z + 3u;
      for (uint32_t lit = 0; lit < leaf_cap; lit++) {
        const int laxis = dda_axis(Lf);
        const uint32_t leaf = (uint32_t)Lf.iz * cell.x * cell.y + (uint32_t)Lf.iy * cell.x + (uint32_t)Lf.ix + cell.w;
        const uint2 range = __ldg(leaf_ranges + leaf);
        for (uint32_t k = range.x; k < range.y; k++) {
          const uint32_t id = __ldg(&leaf_pairs[k].y);
          Vec3 a, b, c; load_tri(tris, id, a, b, c);
          const float t = triangle_intersect(a, b, c, o, d);
          if (t > 0.0f && t < h.t) { h.t = t; h.tri = id; }
        }
        if (h.tri != 0xFFFFFFFFu) return h;                 // contactFound: the first leaf cell with any hit wins
        if (!dda_step(Lf, laxis)) break;
      }
    }
    if (!dda_step(T, axis)) break;
  }
  return none;
}

// ---- corrected walk (ours; SURVEY.md section 8 f4, KNOWN_ISSUES Q21) --------------------------------------------------------------
// The literal walk above enters the grid at t0 = MIN of the slab entries and lets the first leaf cell with any hit end the walk, even
// when that hit lies in a later cell: its contact can differ from the closest hit.  This one is a plain two-level 3D-DDA: entry =
// MAX of the slab entries (clamped to 0), cells visited in ray order, the closest hit so far is kept across cells and the walk ends
// when it lies inside the cell just finished — the true closest hit, the one the BVH returns.  Selected by CLRT_TRACE_FAST on a grid.
struct Walk { float tx, ty, tz, dx, dy, dz; int ix, iy, iz, sx, sy, sz, ex, ey, ez; };
__device__ __forceinline__ void walk_axis(float o, float d, float inv, float b0, float cs, int r, float p, float& tnext, float& dt, int& i, int& step, int& end) {
  i = (int)floorf((p - b0) / cs);
  i = i < 0 ? 0 : (i > r - 1 ? r - 1 : i);
  if (d > 0.0f) { tnext = (b0 + (float)(i + 1) * cs - o) * inv; dt = cs * inv; step = 1; end = r; }
  else if (d < 0.0f) { tnext = (b0 + (float)i * cs - o) * inv; dt = -cs * inv; step = -1; end = -1; }
  else { tnext = RT_FLT_MAX; dt = 0.0f; step = 1; end = r; }
}
__device__ __forceinline__ void walk_setup(Walk& w, Vec3 o, Vec3 d, Vec3 inv, float b0x, float b0y, float b0z, float csx, float csy, float csz,
                                           int rx, int ry, int rz, float t_in) {
  walk_axis(o.x, d.x, inv.x, b0x, csx, rx, o.x + d.x * t_in, w.tx, w.dx, w.ix, w.sx, w.ex);
  walk_axis(o.y, d.y, inv.y, b0y, csy, ry, o.y + d.y * t_in, w.ty, w.dy, w.iy, w.sy, w.ey);
  walk_axis(o.z, d.z, inv.z, b0z, csz, rz, o.z + d.z * t_in, w.tz, w.dz, w.iz, w.sz, w.ez);
}
// leaves the cell through its nearest face; false when that leaves the grid
__device__ __forceinline__ bool walk_step(Walk& w) {
  if (w.tx <= w.ty && w.tx <= w.tz) { w.tx += w.dx; w.ix += w.sx; return w.ix != w.ex; }
  if (w.ty <= w.tz) { w.ty += w.dy; w.iy += w.sy; return w.iy != w.ey; }
  w.tz += w.dz; w.iz += w.sz; return w.iz != w.ez;
}
__device__ __forceinline__ GridHit tlg_traverse_closest(Vec3 o, Vec3 d, const clrt_grid_data& g, const clrt_top_cell* __restrict__ top,
                                                        const uint2* __restrict__ leaf_ranges, const uint2* __restrict__ leaf_pairs,
                                                        const clrt_triangle* __restrict__ tris) {
  GridHit best; best.t = RT_FLT_MAX; best.tri = 0xFFFFFFFFu;
  GridHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
  const float big = 1.0e30f;
  const Vec3 inv = v3(d.x == 0.0f ? copysignf(big, d.x) : 1.0f / d.x, d.y == 0.0f ? copysignf(big, d.y) : 1.0f / d.y, d.z == 0.0f ? copysignf(big, d.z) : 1.0f / d.z);
  const float b0x = g.box.bounds[0].x, b0y = g.box.bounds[0].y, b0z = g.box.bounds[0].z;
  float t_in;
  {
    const float ax = (b0x - o.x) * inv.x, bx = (g.box.bounds[1].x - o.x) * inv.x;
    const float ay = (b0y - o.y) * inv.y, by = (g.box.bounds[1].y - o.y) * inv.y;
    const float az = (b0z - o.z) * inv.z, bz = (g.box.bounds[1].z - o.z) * inv.z;
    const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz)), tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
    if (!(tn <= tf) || tf < 0.0f) return none;
    t_in = fmaxf(tn, 0.0f);
  }
  const float slack = 1.00001f;
  Walk T;
  walk_setup(T, o, d, inv, b0x, b0y, b0z, g.stepX, g.stepY, g.stepZ, (int)g.resX, (int)g.resY, (int)g.resZ, t_in);
  float t_cur = t_in;
  const uint32_t top_cap = g.resX + g.resY + g.resZ + 3u;
  for (uint32_t it = 0; it < top_cap; it++) {
    const uint4 cell = __ldg(reinterpret_cast<const uint4*>(top) + ((uint32_t)T.iz * g.resX * g.resY + (uint32_t)T.iy * g.resX + (uint32_t)T.ix));
    const float t_exit = fminf(T.tx, fminf(T.ty, T.tz));
    if (!(cell.x == 0 || cell.y == 0 || cell.z == 0)) {
      const float c0x = b0x + (float)T.ix * g.stepX, c0y = b0y + (float)T.iy * g.stepY, c0z = b0z + (float)T.iz * g.stepZ;
      Walk Lf;
      walk_setup(Lf, o, d, inv, c0x, c0y, c0z, g.stepX / (float)cell.x, g.stepY / (float)cell.y, g.stepZ / (float)cell.z,
                 (int)cell.x, (int)cell.y, (int)cell.z, t_cur);
      const uint32_t leaf_cap = cell.x + cell.y + cell.z + 3u;
      for (uint32_t lit = 0; lit < leaf_cap; lit++) {
        const uint32_t leaf = (uint32_t)Lf.iz * cell.x * cell.y + (uint32_t)Lf.iy * cell.x + (uint32_t)Lf.ix + cell.w;
        const uint2 range = __ldg(leaf_ranges + leaf);
        for (uint32_t k = range.x; k < range.y; k++) {
          const uint32_t id = __ldg(&leaf_pairs[k].y);
          Vec3 a, b, c; load_tri(tris, id, a, b, c);
          const float t = triangle_intersect(a, b, c, o, d);
          // bit-equal t: the reference's traversal keeps the triangle that comes first in its sorted (Morton code, index) order
          if (t > 0.0f && (t < best.t || (t == best.t && reference_order_before(tris, id, best.tri, g.box)))) { best.t = t; best.tri = id; }
        }
        const float l_exit = fminf(Lf.tx, fminf(Lf.ty, Lf.tz));
        if (best.t <= l_exit * slack) return best;            // the closest hit lies in a cell already finished
        if (l_exit >= t_exit || !walk_step(Lf)) break;        // the ray left the top cell
      }
    }
    if (best.t <= t_exit * slack) return best;
    t_cur = t_exit;
    if (!walk_step(T)) break;
  }
  return best.tri != 0xFFFFFFFFu ? best : none;
}

__device__ __forceinline__ void tlg_write_contact(clrt_contact* __restrict__ out, uint32_t slot, uint32_t pixelIndex, const GridHit& h,
                                                  const clrt_triangle* __restrict__ tris) {
  uint4 head = make_uint4(pixelIndex, 0u, 0xFFFFFFFFu, 0u);
  float4 nd = make_float4(0.0f, 0.0f, 0.0f, 0.0f);                 // NO_CONTACT (CLStructs.h:124)
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

constexpr int TLG_TRACE_THREADS = 128;

// generateContactsKernel (.cl:228-246): persistent warps over 8x4-pixel packets (same tiling as the BVH path).
template <bool CORRECTED>
__global__ void __launch_bounds__(TLG_TRACE_THREADS)
tlg_trace_primary_kernel(const clrt_camera cam, const clrt_grid_data g, const clrt_top_cell* __restrict__ top,
                         const uint2* __restrict__ leaf_ranges, const uint2* __restrict__ leaf_pairs, const clrt_triangle* __restrict__ tris,
                         clrt_contact* __restrict__ out, uint32_t rank, uint32_t nranks, int raster, uint32_t* __restrict__ counter) {
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
      const GridHit h = CORRECTED ? tlg_traverse_closest(o, d, g, top, leaf_ranges, leaf_pairs, tris) : tlg_traverse(o, d, g, top, leaf_ranges, leaf_pairs, tris);
      tlg_write_contact(out, raster ? pixel : lt * CLRT_TILE_PIXELS + inTile, pixel, h, tris);
    } else if (!raster) {
      GridHit h; h.t = 0.0f; h.tri = 0xFFFFFFFFu;
      tlg_write_contact(out, lt * CLRT_TILE_PIXELS + inTile, 0xFFFFFFFFu, h, tris);
    }
  }
}

// generateContacts2Kernel (.cl:251-267)
template <bool CORRECTED>
__global__ void __launch_bounds__(TLG_TRACE_THREADS)
tlg_trace_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const clrt_grid_data g, const clrt_top_cell* __restrict__ top,
                      const uint2* __restrict__ leaf_ranges, const uint2* __restrict__ leaf_pairs, const clrt_triangle* __restrict__ tris,
                      clrt_contact* __restrict__ out, uint32_t* __restrict__ counter) {
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
      GridHit h; h.t = 0.0f; h.tri = 0xFFFFFFFFu;
      if (!(rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f))
        h = CORRECTED ? tlg_traverse_closest(v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), g, top, leaf_ranges, leaf_pairs, tris)
                      : tlg_traverse(v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z), g, top, leaf_ranges, leaf_pairs, tris);
      tlg_write_contact(out, i, i, h, tris);
    }
  }
}

// ---- the corrected walk on persistent warps with lane refill ------------------------------------------------------------------------
// tlg_traverse_closest above runs one ray per thread to the end: rays end after one top cell or after fifty, so a warp ran with 8-9 of
// 32 lanes (ncu, profiles/r02_all_kernels.md).  Here the SAME walk (same arithmetic, same order of operations per ray, hence the same
// hits) is resumable between TOP CELLS: one step = one whole top cell of the ray (its leaf walk and triangle loops as compact nested
// loops), lanes meet again between top cells, and complete rays wait until enough lanes wait, then write their contacts and are
// refilled together from an atomic counter (the LBVH kernels' scheme, trace.cu run_jobs).  A finer state machine — separate warp-wide
// phases for "advance to the next leaf cell" and "test one triangle" — was measured slower than no refill at all (6.3 ms against
// 4.76 ms per Soup-1M frame; this scheme: 4.10 ms): its 100 registers and branchy steps cost more than the lanes it recovered
// (profiles/r02_experiments.md section 7).
#ifndef CLRT_TLG_REFILL
#define CLRT_TLG_REFILL 1            /* 0: the corrected walk runs one ray per thread (tlg_trace_*_kernel<true>) */
#endif
#ifndef CLRT_TLG_WINDOW
#define CLRT_TLG_WINDOW 1            /* top cells per lane between refill checks */
#endif
#ifndef CLRT_TLG_REFILL_MIN
#define CLRT_TLG_REFILL_MIN 8
#endif
#ifndef CLRT_TLG_MIN_CTAS
#define CLRT_TLG_MIN_CTAS 7
#endif
// The walk state without the step / end members (both follow from the sign of d and the resolution): six registers fewer per walk in the
// lane-refill kernels, where the top-level walk lives across steps.
struct WalkC { float tx, ty, tz, dx, dy, dz; int ix, iy, iz; };
__device__ __forceinline__ void walkc_axis(float o, float d, float inv, float b0, float cs, int r, float p, float& tnext, float& dt, int& i) {
  i = (int)floorf((p - b0) / cs);
  i = i < 0 ? 0 : (i > r - 1 ? r - 1 : i);
  if (d > 0.0f) { tnext = (b0 + (float)(i + 1) * cs - o) * inv; dt = cs * inv; }
  else if (d < 0.0f) { tnext = (b0 + (float)i * cs - o) * inv; dt = -cs * inv; }
  else { tnext = RT_FLT_MAX; dt = 0.0f; }
}
__device__ __forceinline__ void walkc_setup(WalkC& w, Vec3 o, Vec3 d, Vec3 inv, float b0x, float b0y, float b0z, float csx, float csy, float csz,
                                            int rx, int ry, int rz, float t_in) {
  walkc_axis(o.x, d.x, inv.x, b0x, csx, rx, o.x + d.x * t_in, w.tx, w.dx, w.ix);
  walkc_axis(o.y, d.y, inv.y, b0y, csy, ry, o.y + d.y * t_in, w.ty, w.dy, w.iy);
  walkc_axis(o.z, d.z, inv.z, b0z, csz, rz, o.z + d.z * t_in, w.tz, w.dz, w.iz);
}
// leaves the cell through its nearest face; false when that leaves the grid (an axis with d == 0 is never chosen: its tnext is FLT_MAX)
__device__ __forceinline__ bool walkc_step(WalkC& w, Vec3 d, int rx, int ry, int rz) {
  if (w.tx <= w.ty && w.tx <= w.tz) { w.tx += w.dx; if (d.x < 0.0f) { w.ix--; return w.ix != -1; } w.ix++; return w.ix != rx; }
  if (w.ty <= w.tz) { w.ty += w.dy; if (d.y < 0.0f) { w.iy--; return w.iy != -1; } w.iy++; return w.iy != ry; }
  w.tz += w.dz; if (d.z < 0.0f) { w.iz--; return w.iz != -1; } w.iz++; return w.iz != rz;
}
struct GLane {
  Vec3 o, d, inv;
  GridHit best;
  WalkC T;
  float t_cur;
  uint32_t it;
  bool done;
  uint32_t slot, pixel;
};
struct GridView {
  clrt_grid_data g; const clrt_top_cell* top; const uint2* leaf_ranges; const uint2* leaf_pairs; const clrt_triangle* tris;
};

// true when the ray enters the grid; false: the contact is "none" already
__device__ __forceinline__ bool glane_begin(GLane& L, const GridView& V, Vec3 o, Vec3 d) {
  const clrt_grid_data& g = V.g;
  L.o = o; L.d = d;
  L.best.t = RT_FLT_MAX; L.best.tri = 0xFFFFFFFFu;
  const float big = 1.0e30f;
  L.inv = v3(d.x == 0.0f ? copysignf(big, d.x) : 1.0f / d.x, d.y == 0.0f ? copysignf(big, d.y) : 1.0f / d.y, d.z == 0.0f ? copysignf(big, d.z) : 1.0f / d.z);
  const float b0x = g.box.bounds[0].x, b0y = g.box.bounds[0].y, b0z = g.box.bounds[0].z;
  const float ax = (b0x - o.x) * L.inv.x, bx = (g.box.bounds[1].x - o.x) * L.inv.x;
  const float ay = (b0y - o.y) * L.inv.y, by = (g.box.bounds[1].y - o.y) * L.inv.y;
  const float az = (b0z - o.z) * L.inv.z, bz = (g.box.bounds[1].z - o.z) * L.inv.z;
  const float tn = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fminf(az, bz)), tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz));
  if (!(tn <= tf) || tf < 0.0f) { L.done = true; return false; }
  const float t_in = fmaxf(tn, 0.0f);
  walkc_setup(L.T, o, d, L.inv, b0x, b0y, b0z, g.stepX, g.stepY, g.stepZ, (int)g.resX, (int)g.resY, (int)g.resZ, t_in);
  L.t_cur = t_in; L.it = 0; L.done = false;
  return true;
}
// One step: one whole top cell of the ray (its leaf walk and triangle loops exactly as in tlg_traverse_closest).
__device__ __forceinline__ void glane_top_cell(GLane& L, const GridView& V) {
  const clrt_grid_data& g = V.g;
  const float slack = 1.00001f;
  const uint4 cell = __ldg(reinterpret_cast<const uint4*>(V.top) + ((uint32_t)L.T.iz * g.resX * g.resY + (uint32_t)L.T.iy * g.resX + (uint32_t)L.T.ix));
  const float t_exit = fminf(L.T.tx, fminf(L.T.ty, L.T.tz));
  if (!(cell.x == 0 || cell.y == 0 || cell.z == 0)) {
    const float c0x = g.box.bounds[0].x + (float)L.T.ix * g.stepX, c0y = g.box.bounds[0].y + (float)L.T.iy * g.stepY, c0z = g.box.bounds[0].z + (float)L.T.iz * g.stepZ;
    WalkC Lf;
    walkc_setup(Lf, L.o, L.d, L.inv, c0x, c0y, c0z, g.stepX / (float)cell.x, g.stepY / (float)cell.y, g.stepZ / (float)cell.z,
               (int)cell.x, (int)cell.y, (int)cell.z, L.t_cur);
    const uint32_t leaf_cap = cell.x + cell.y + cell.z + 3u;
    for (uint32_t lit = 0; lit < leaf_cap; lit++) {
      const uint32_t leaf = (uint32_t)Lf.iz * cell.x * cell.y + (uint32_t)Lf.iy * cell.x + (uint32_t)Lf.ix + cell.w;
      const uint2 range = __ldg(V.leaf_ranges + leaf);
      for (uint32_t k = range.x; k < range.y; k++) {
        const uint32_t id = __ldg(&V.leaf_pairs[k].y);
        Vec3 a, b, c; load_tri(V.tris, id, a, b, c);
        const float t = triangle_intersect(a, b, c, L.o, L.d);
        if (t > 0.0f && (t < L.best.t || (t == L.best.t && reference_order_before(V.tris, id, L.best.tri, g.box)))) { L.best.t = t; L.best.tri = id; }
      }
      const float l_exit = fminf(Lf.tx, fminf(Lf.ty, Lf.tz));
      if (L.best.t <= l_exit * slack) { L.done = true; return; }
      if (l_exit >= t_exit || !walkc_step(Lf, L.d, (int)cell.x, (int)cell.y, (int)cell.z)) break;
    }
  }
  if (L.best.t <= t_exit * slack) { L.done = true; return; }
  L.t_cur = t_exit;
  L.it++;
  if (!walkc_step(L.T, L.d, (int)g.resX, (int)g.resY, (int)g.resZ) || L.it >= g.resX + g.resY + g.resZ + 3u) L.done = true;
}

// Job: start(unit, L) -> the lane is busy; the contact of a ray that never enters the grid (or of a slot outside the image) is written at once
template <class Job>
__device__ __forceinline__ void tlg_run_jobs(Job& job, const GridView& V, uint32_t total, uint32_t* __restrict__ counter, clrt_contact* __restrict__ out) {
  const unsigned FULL = 0xffffffffu;
  const int lane = threadIdx.x & 31;
  const unsigned lt_mask = (1u << lane) - 1u;
  GLane L; L.done = true;
  bool busy = false, exhausted = false;
  uint32_t pool_next = 0, pool_end = 0;
  while (true) {
    const bool fin = busy && L.done;
    const unsigned waiting = __ballot_sync(FULL, !busy || fin);
    if (__popc(waiting) >= CLRT_TLG_REFILL_MIN || !__any_sync(FULL, busy && !fin)) {
      if (fin) {
        GridHit h = L.best; if (h.tri == 0xFFFFFFFFu) h.t = 0.0f;
        tlg_write_contact(out, L.slot, L.pixel, h, V.tris);
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
      if (idle == FULL) break;                                   // nothing left anywhere
    }
#pragma unroll 1
    for (int w = 0; w < CLRT_TLG_WINDOW; w++) if (busy && !L.done) glane_top_cell(L, V);
  }
}

struct GridPrimaryJob {
  clrt_camera cam; TileGrid tg; uint32_t rank, nranks; int raster;
  __device__ __forceinline__ bool start(uint32_t unit, GLane& L, const GridView& V, clrt_contact* __restrict__ out) const {
    const uint32_t pk = unit >> 5, ln = unit & 31u;
    const uint32_t lt = pk >> 3, sub = pk & 7u;
    const uint32_t gt = rank + lt * nranks;
    const uint32_t tx = (gt % tg.tilesX) * CLRT_TILE_W + (sub & 3u) * 8u + (ln & 7u);
    const uint32_t ty = (gt / tg.tilesX) * CLRT_TILE_H + (sub >> 2) * 4u + (ln >> 3);
    const uint32_t inTile = ((sub >> 2) * 4u + (ln >> 3)) * CLRT_TILE_W + (sub & 3u) * 8u + (ln & 7u);
    const bool inside = tx < cam.resX && ty < cam.resY;
    const uint32_t pixel = ty * cam.resX + tx;
    GridHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
    if (!inside) { if (!raster) tlg_write_contact(out, lt * CLRT_TILE_PIXELS + inTile, 0xFFFFFFFFu, none, V.tris); return false; }
    L.slot = raster ? pixel : lt * CLRT_TILE_PIXELS + inTile; L.pixel = pixel;
    Vec3 o, d; generate_ray(cam, pixel, o, d);
    if (!glane_begin(L, V, o, d)) { tlg_write_contact(out, L.slot, pixel, none, V.tris); return false; }
    return true;
  }
};
struct GridRaysJob {
  const clrt_ray* rays;
  __device__ __forceinline__ bool start(uint32_t i, GLane& L, const GridView& V, clrt_contact* __restrict__ out) const {
    const float4* p = reinterpret_cast<const float4*>(rays + i);
    const float4 ro = __ldcs(p + 1), rd = __ldcs(p + 2);
    GridHit none; none.t = 0.0f; none.tri = 0xFFFFFFFFu;
    L.slot = i; L.pixel = i;
    if ((rd.x == 0.0f && rd.y == 0.0f && rd.z == 0.0f) || !glane_begin(L, V, v3(ro.x, ro.y, ro.z), v3(rd.x, rd.y, rd.z))) {
      tlg_write_contact(out, i, i, none, V.tris); return false;
    }
    return true;
  }
};

__global__ void __launch_bounds__(TLG_TRACE_THREADS, CLRT_TLG_MIN_CTAS)
tlg_walk_primary_kernel(const clrt_camera cam, const GridView V, clrt_contact* __restrict__ out, uint32_t rank, uint32_t nranks, int raster,
                        uint32_t* __restrict__ counter) {
  GridPrimaryJob job{cam, make_tile_grid(cam.resX, cam.resY), rank, nranks, raster};
  const uint32_t total = stripe_tile_count(job.tg.tiles, rank, nranks) * 8u * 32u;
  tlg_run_jobs(job, V, total, counter, out);
}
__global__ void __launch_bounds__(TLG_TRACE_THREADS, CLRT_TLG_MIN_CTAS)
tlg_walk_rays_kernel(const clrt_ray* __restrict__ rays, uint32_t count, const GridView V, clrt_contact* __restrict__ out, uint32_t* __restrict__ counter) {
  GridRaysJob job{rays};
  tlg_run_jobs(job, V, count, counter, out);
}

} // namespace

// ---- manager ---------------------------------------------------------------------------------------------------------
static int bits_for(uint32_t max_value) { int b = 1; while (b < 32 && (max_value >> b)) b++; return b; }

static int read_u32(clrt_ctx* ctx, const uint32_t* d, uint32_t* out) {
  void* pin = nullptr;
  CLRT_TRY(clrt_ctx_pinned(ctx, 64, &pin));
  CLRT_CUDA(cudaMemcpyAsync(pin, d, 4, cudaMemcpyDeviceToHost, ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(ctx->stream));
  *out = *(uint32_t*)pin;
  return 0;
}
static int grow_sync(clrt_ctx* ctx, void** p, size_t* cap, size_t bytes) {
  if (bytes > *cap || !*p) { CLRT_CUDA(cudaStreamSynchronize(ctx->stream)); CLRT_TRY(clrt_grow(p, cap, bytes)); }
  return 0;
}

// getResolution (TwoLevelGridManager.cpp:566-578): float arithmetic, pow in double rounded once (Q20), truncation to uint.
static void tlg_resolution(const clrt_accel* a, float& fx, float& fy, float& fz) {
  const clrt_aabb& box = a->scene->header.modelsBoundingBox;
  const float deltaX = box.bounds[1].x - box.bounds[0].x, deltaY = box.bounds[1].y - box.bounds[0].y, deltaZ = box.bounds[1].z - box.bounds[0].z;
  const float volume = deltaX * deltaY * deltaZ;
  const float prims = (float)a->scene->header.totalNumberOfTriangles;
  const float av = (float)pow((double)(a->top_density * prims / volume), (double)ONE_THIRD);
  fx = deltaX * av; fy = deltaY * av; fz = deltaZ * av;
}
// setTopLevelDensity / setLeafDensity (TwoLevelGridManager.h:112-115): plain members read by calculateGridData at the next initializeFrame.
extern "C" int clrt_tlg_set_densities(clrt_accel* a, float top_density, float leaf_density) {
  if (!a || a->kind != ACCEL_TLG) CLRT_FAIL("clrt_tlg_set_densities: not a Two-Level Grid");
  if (!(top_density > 0.0f) || !(leaf_density > 0.0f)) CLRT_FAIL("clrt_tlg_set_densities: densities must be positive");
  a->top_density = top_density; a->leaf_density = leaf_density;
  return 0;
}
// getResolution() (TwoLevelGridManager.h:129): the top-level resolution the current densities give, truncated like the reference's
// float -> uint conversion (0 where the reference's value is below one or not a number).
extern "C" int clrt_tlg_resolution(clrt_accel* a, uint32_t* out_xyz) {
  if (!a || !out_xyz || a->kind != ACCEL_TLG) CLRT_FAIL("clrt_tlg_resolution: not a Two-Level Grid / NULL argument");
  float f[3]; tlg_resolution(a, f[0], f[1], f[2]);
  for (int i = 0; i < 3; i++) out_xyz[i] = (f[i] >= 1.0f && f[i] < 4294967040.0f) ? (uint32_t)f[i] : 0u;
  return 0;
}

// initializeFrame (TwoLevelGridManager.cpp:160-201): calculateGridData (:584-600) on the host, allocate + zero.
int tlg_init_frame(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  if (!a->tlg) { a->tlg = new TlgState(); }
  TlgState* t = a->tlg;
  t->built = false; a->built = false;
  const clrt_scene* sc = a->scene;
  const uint32_t n = sc->n_tris;
  a->n = n;
  clrt_grid_data g; memset(&g, 0, sizeof g);
  g.box = sc->header.modelsBoundingBox;
  {
    float fx, fy, fz; tlg_resolution(a, fx, fy, fz);
    if (!(fx >= 1.0f && fy >= 1.0f && fz >= 1.0f) || !(fx < 2048.0f && fy < 2048.0f && fz < 2048.0f) || n == 0)
      CLRT_FAIL("Two-Level Grid: degenerate top-level resolution (%g x %g x %g) for %u triangles - the reference divides by zero here", fx, fy, fz, n);
    g.resX = (uint32_t)fx; g.resY = (uint32_t)fy; g.resZ = (uint32_t)fz;
    g.stepX = (g.box.bounds[1].x - g.box.bounds[0].x) / (float)g.resX;
    g.stepY = (g.box.bounds[1].y - g.box.bounds[0].y) / (float)g.resY;
    g.stepZ = (g.box.bounds[1].z - g.box.bounds[0].z) / (float)g.resZ;
    g.leafDensity = a->leaf_density;
  }
  t->grid = g;
  t->cells = g.resX * g.resY * g.resZ;
  const size_t words = (size_t)(n > t->cells ? n : t->cells);
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_counters, &t->counters_cap, words * 4));
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_scan, &t->scan_cap, words * 4));
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_top_ranges, &t->top_ranges_cap, (size_t)t->cells * 8));
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_top_cells, &t->top_cells_cap, (size_t)t->cells * 16));
  if (!t->d_real) CLRT_CUDA(cudaMalloc(&t->d_real, 64));
  CLRT_CUDA(cudaMemsetAsync(t->d_top_ranges, 0, (size_t)t->cells * 8, ctx->stream));
  CLRT_CUDA(cudaMemsetAsync(t->d_top_cells, 0, (size_t)t->cells * 16, ctx->stream));
  CLRT_CUDA(cudaMemsetAsync(t->d_real, 0, 64, ctx->stream));
  return 0;
}

int tlg_construct(clrt_accel* a) {
  clrt_ctx* ctx = a->ctx;
  TlgState* t = a->tlg;
  if (!t || !t->d_counters) CLRT_FAIL("Two-Level Grid construct: initializeFrame was not called");
  if (a->n != a->scene->n_tris) CLRT_FAIL("Two-Level Grid construct: scene changed since initializeFrame");
  cudaStream_t s = ctx->stream;
  const uint32_t n = a->n;
  const clrt_grid_data g = t->grid;
  const clrt_triangle* tris = a->scene->d_tris;
  const bool tm = a->timing;
  int ev = 0;
  auto mark = [&]() -> int { if (tm) CLRT_CUDA(cudaEventRecord(a->ev[ev++], s)); return 0; };
  CLRT_TRY(mark());                                                                     // ev0
  // 1-2. count + scan + pair count
  tlg_count_top_kernel<<<(n + 255) / 256, 256, 0, s>>>(tris, n, g, t->d_counters);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(scan_inclusive(ctx, t->d_counters, t->d_scan, n));
  CLRT_TRY(read_u32(ctx, t->d_scan + (n - 1), &t->top_pairs));
  if (t->top_pairs == 0 || t->top_pairs >= (1u << 30)) CLRT_FAIL("Two-Level Grid: %u top-level pairs", t->top_pairs);
  CLRT_TRY(mark());                                                                     // ev1: count
  // 3-5. scatter, sort, ranges
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_top_pairs, &t->top_pairs_cap, (size_t)t->top_pairs * 8));
  tlg_scatter_top_kernel<<<(n + 255) / 256, 256, 0, s>>>(tris, n, g, t->d_scan, t->d_counters, t->top_pairs, (uint2*)t->d_top_pairs);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(sort_pairs_radix(ctx, t->d_top_pairs, t->top_pairs, bits_for(t->cells - 1)));
  extract_ranges_kernel<<<(t->top_pairs + 255) / 256, 256, 0, s>>>((const uint2*)t->d_top_pairs, t->top_pairs, (uint2*)t->d_top_ranges, nullptr);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(mark());                                                                     // ev2: top pairs
  // 6-7. leaf resolutions, leaf-cell offsets
  tlg_leaf_res_kernel<<<(t->cells + 255) / 256, 256, 0, s>>>((const uint2*)t->d_top_ranges, t->cells, g, t->d_top_cells, t->d_counters);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(scan_inclusive(ctx, t->d_counters, t->d_scan, t->cells));
  tlg_first_leaf_kernel<<<(t->cells + 255) / 256, 256, 0, s>>>(t->d_top_cells, t->d_scan, t->cells);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(read_u32(ctx, t->d_scan + (t->cells - 1), &t->leaf_cells));
  CLRT_TRY(mark());                                                                     // ev3: leaf cells
  // 8-9. conservative leaf-pair counts
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_counters, &t->counters_cap, (size_t)t->top_pairs * 4));
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_scan, &t->scan_cap, (size_t)t->top_pairs * 4));
  tlg_count_leaf_kernel<<<(t->top_pairs + 127) / 128, 128, 0, s>>>(tris, (const uint2*)t->d_top_pairs, t->top_pairs, g, t->d_top_cells, t->d_counters);
  CLRT_CHECK_LAUNCH();
  CLRT_TRY(scan_inclusive(ctx, t->d_counters, t->d_scan, t->top_pairs));
  CLRT_TRY(read_u32(ctx, t->d_scan + (t->top_pairs - 1), &t->leaf_pairs_cap));
  if (t->leaf_pairs_cap >= (1u << 30)) CLRT_FAIL("Two-Level Grid: %u leaf pairs", t->leaf_pairs_cap);
  CLRT_TRY(mark());                                                                     // ev4: leaf count
  // 10-12. exact scatter, sort (gaps = UINT_MAX sort last), ranges
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_leaf_pairs, &t->leaf_pairs_capacity, (size_t)(t->leaf_pairs_cap ? t->leaf_pairs_cap : 1) * 8));
  CLRT_TRY(grow_sync(ctx, (void**)&t->d_leaf_ranges, &t->leaf_ranges_cap, (size_t)(t->leaf_cells ? t->leaf_cells : 1) * 8));
  CLRT_CUDA(cudaMemsetAsync(t->d_leaf_pairs, 0xFF, (size_t)t->leaf_pairs_cap * 8, s));
  CLRT_CUDA(cudaMemsetAsync(t->d_leaf_ranges, 0, (size_t)t->leaf_cells * 8, s));
  CLRT_CUDA(cudaMemsetAsync(t->d_real, 0, 8, s));                                      // [0] real leaf pairs, [1] pairs beyond their budget
  t->leaf_pairs = 0;
  if (t->leaf_pairs_cap) {
    tlg_scatter_leaf_kernel<<<(t->top_pairs + 127) / 128, 128, 0, s>>>(tris, (const uint2*)t->d_top_pairs, t->top_pairs, g, t->d_top_cells,
                                                                       t->d_scan, t->d_counters, t->leaf_pairs_cap, (uint2*)t->d_leaf_pairs, t->d_real + 1);
    CLRT_CHECK_LAUNCH();
    CLRT_TRY(sort_pairs_radix(ctx, t->d_leaf_pairs, t->leaf_pairs_cap, 32));
    extract_ranges_kernel<<<(t->leaf_pairs_cap + 255) / 256, 256, 0, s>>>((const uint2*)t->d_leaf_pairs, t->leaf_pairs_cap, (uint2*)t->d_leaf_ranges, t->d_real);
    CLRT_CHECK_LAUNCH();
    CLRT_TRY(read_u32(ctx, t->d_real, &t->leaf_pairs));
    uint32_t over = 0;
    CLRT_TRY(read_u32(ctx, t->d_real + 1, &over));
    if (over) CLRT_FAIL("Two-Level Grid: the exact leaf scatter found %u pairs the conservative count did not budget for", over);
  }
  CLRT_TRY(mark());                                                                     // ev5: leaf pairs
  if (tm) {
    CLRT_CUDA(cudaEventSynchronize(a->ev[5]));
    for (int i = 0; i < 5; i++) CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[8 + i], a->ev[i], a->ev[i + 1]));
    CLRT_CUDA(cudaEventElapsedTime(&a->stage_ms[CLRT_STAGE_TOTAL], a->ev[0], a->ev[5]));
  }
  t->built = true;
  a->built = true;
  return 0;
}

template <typename K>
static int tlg_grid(clrt_ctx* ctx, K kernel, uint32_t packets, unsigned* grid) {
  int per_sm = 0;
  CLRT_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, TLG_TRACE_THREADS, 0));
  if (per_sm < 1) per_sm = 1;
  *grid = min((unsigned)(ctx->sm_count * per_sm), (packets + 3) / 4);
  return 0;
}

int tlg_trace_primary(clrt_accel* a, const clrt_camera& cam, clrt_contact* out, uint32_t rank, uint32_t nranks) {
  clrt_ctx* ctx = a->ctx; TlgState* t = a->tlg;
  if (!t || !t->built) CLRT_FAIL("generateContacts: the Two-Level Grid was not constructed");
  const bool raster = (nranks == 0);
  if (raster) { rank = 0; nranks = 1; }
  const TileGrid tg = make_tile_grid(cam.resX, cam.resY);
  const uint32_t packets = stripe_tile_count(tg.tiles, rank, nranks) * 8u;
  if (!packets) return 0;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  unsigned grid = 0;
  if (a->trace_mode == CLRT_TRACE_FAST && CLRT_TLG_REFILL) {   // the corrected closest-hit walk, lanes refilled
    const GridView V{t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges, (const uint2*)t->d_leaf_pairs, a->scene->d_tris};
    CLRT_TRY(tlg_grid(ctx, tlg_walk_primary_kernel, packets, &grid));
    tlg_walk_primary_kernel<<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(cam, V, out, rank, nranks, raster, a->d_tile_counter);
  } else if (a->trace_mode == CLRT_TRACE_FAST) {               // the corrected closest-hit walk, one ray per thread
    CLRT_TRY(tlg_grid(ctx, tlg_trace_primary_kernel<true>, packets, &grid));
    tlg_trace_primary_kernel<true><<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(cam, t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges,
        (const uint2*)t->d_leaf_pairs, a->scene->d_tris, out, rank, nranks, raster, a->d_tile_counter);
  } else {                                                     // the reference's walk, literally
    CLRT_TRY(tlg_grid(ctx, tlg_trace_primary_kernel<false>, packets, &grid));
    tlg_trace_primary_kernel<false><<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(cam, t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges,
        (const uint2*)t->d_leaf_pairs, a->scene->d_tris, out, rank, nranks, raster, a->d_tile_counter);
  }
  CLRT_CHECK_LAUNCH();
  return 0;
}

int tlg_trace_rays(clrt_accel* a, const clrt_ray* rays, clrt_contact* out, uint32_t count) {
  clrt_ctx* ctx = a->ctx; TlgState* t = a->tlg;
  if (!t || !t->built) CLRT_FAIL("generateContacts: the Two-Level Grid was not constructed");
  if (!count) return 0;
  const uint32_t packets = (count + 31u) / 32u;
  CLRT_CUDA(cudaMemsetAsync(a->d_tile_counter, 0, 4, ctx->stream));
  unsigned grid = 0;
  if (a->trace_mode == CLRT_TRACE_FAST && CLRT_TLG_REFILL) {
    const GridView V{t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges, (const uint2*)t->d_leaf_pairs, a->scene->d_tris};
    CLRT_TRY(tlg_grid(ctx, tlg_walk_rays_kernel, packets, &grid));
    tlg_walk_rays_kernel<<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(rays, count, V, out, a->d_tile_counter);
  } else if (a->trace_mode == CLRT_TRACE_FAST) {
    CLRT_TRY(tlg_grid(ctx, tlg_trace_rays_kernel<true>, packets, &grid));
    tlg_trace_rays_kernel<true><<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(rays, count, t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges,
        (const uint2*)t->d_leaf_pairs, a->scene->d_tris, out, a->d_tile_counter);
  } else {
    CLRT_TRY(tlg_grid(ctx, tlg_trace_rays_kernel<false>, packets, &grid));
    tlg_trace_rays_kernel<false><<<grid, TLG_TRACE_THREADS, 0, ctx->stream>>>(rays, count, t->grid, t->d_top_cells, (const uint2*)t->d_leaf_ranges,
        (const uint2*)t->d_leaf_pairs, a->scene->d_tris, out, a->d_tile_counter);
  }
  CLRT_CHECK_LAUNCH();
  return 0;
}

void tlg_destroy(clrt_accel* a) {
  TlgState* t = a->tlg;
  if (!t) return;
  void* ptrs[] = {t->d_counters, t->d_scan, t->d_top_ranges, t->d_top_cells, t->d_top_pairs, t->d_leaf_pairs, t->d_leaf_ranges, t->d_real};
  for (void* p : ptrs) if (p) cudaFree(p);
  delete t; a->tlg = nullptr;
}

extern "C" int clrt_tlg_sizes(clrt_accel* a, uint32_t* top_cells, uint32_t* top_pairs, uint32_t* leaf_cells, uint32_t* leaf_pairs) {
  if (!a || a->kind != ACCEL_TLG || !a->tlg || !a->tlg->built) CLRT_FAIL("clrt_tlg_sizes: not a constructed Two-Level Grid");
  const TlgState* t = a->tlg;
  if (top_cells) *top_cells = t->cells;
  if (top_pairs) *top_pairs = t->top_pairs;
  if (leaf_cells) *leaf_cells = t->leaf_cells;
  if (leaf_pairs) *leaf_pairs = t->leaf_pairs;
  return 0;
}

extern "C" int clrt_tlg_export(clrt_accel* a, clrt_grid_data* h_grid, void* h_top_cells, void* h_top_ranges, void* h_top_pairs,
                               void* h_leaf_ranges, void* h_leaf_pairs) {
  if (!a || a->kind != ACCEL_TLG || !a->tlg || !a->tlg->built) CLRT_FAIL("clrt_tlg_export: not a constructed Two-Level Grid");
  const TlgState* t = a->tlg; cudaStream_t s = a->ctx->stream;
  CLRT_CUDA(cudaSetDevice(a->ctx->device));
  if (h_grid) *h_grid = t->grid;
  if (h_top_cells) CLRT_CUDA(cudaMemcpyAsync(h_top_cells, t->d_top_cells, (size_t)t->cells * 16, cudaMemcpyDeviceToHost, s));
  if (h_top_ranges) CLRT_CUDA(cudaMemcpyAsync(h_top_ranges, t->d_top_ranges, (size_t)t->cells * 8, cudaMemcpyDeviceToHost, s));
  if (h_top_pairs) CLRT_CUDA(cudaMemcpyAsync(h_top_pairs, t->d_top_pairs, (size_t)t->top_pairs * 8, cudaMemcpyDeviceToHost, s));
  if (h_leaf_ranges && t->leaf_cells) CLRT_CUDA(cudaMemcpyAsync(h_leaf_ranges, t->d_leaf_ranges, (size_t)t->leaf_cells * 8, cudaMemcpyDeviceToHost, s));
  if (h_leaf_pairs && t->leaf_pairs) CLRT_CUDA(cudaMemcpyAsync(h_leaf_pairs, t->d_leaf_pairs, (size_t)t->leaf_pairs * 8, cudaMemcpyDeviceToHost, s));
  CLRT_CUDA(cudaStreamSynchronize(s));
  return 0;
}
