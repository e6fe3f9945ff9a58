// oracle/clrt_oracle.cpp — TEST INFRASTRUCTURE.  CPU restatement of the reference's algorithms for the hot path.
//
// Nothing here is part of the product: only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs load this library, and only as the checker (or as the timed CPU baseline).  The product (libclrt_cuda.so) never
// links or calls it and has no CPU fallback.
//
// Pinning: the reference ships no golden vectors.  This restatement is pinned against the reference's
// OWN headers compiled in place (oracle/ref_bridge.cpp -> oracle/_ref/libclrt_ref.so) by tests/test_oracle_pin.py, and
// through fixtures generated from that library (tests/golden/, script tests/golden/make_golden.py).
//
// R = /root/reference/OpenCLRayTracer.  Each function cites the reference lines it restates.  Arithmetic notes:
// built with -ffp-contract=off (no FMA), `rmin/rmax` are the reference's host macros (compare-select,
// R/Include/CLData/CLPortability.h:147-148), normalisation is the exact 1/sqrt form (:151-159), `sign` has OpenCL
// semantics (the host macro at :282 is a precedence bug).
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <limits.h>
#include <algorithm>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif
#include "../include/clrt_types.h"

namespace {

inline float rmin(float a, float b) { return (a < b) ? a : b; }
inline float rmax(float a, float b) { return (a > b) ? a : b; }
struct V3 { float x, y, z; };
inline V3 v3(float x, float y, float z) { V3 r = {x, y, z}; return r; }
inline V3 operator-(V3 a, V3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator+(V3 a, V3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator*(V3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
inline V3 cross(V3 a, V3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }  // CLPortability.h:267-274
inline float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }                                          // :276-279
inline V3 normalize(V3 a) { const float l = 1.0f / sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); return v3(a.x * l, a.y * l, a.z * l); }  // :151-159
inline float length3(V3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }                                     // LENGTH3 :279
inline V3 xyz(const clrt_float4& f) { return v3(f.x, f.y, f.z); }
inline uint32_t clz32(uint32_t x) { return x ? (uint32_t)__builtin_clz(x) : 32u; }                                  // software_clz :163-174

// ---- scene access: one header walk at load instead of per fetch (SceneBufferParser.h:175-195,253-293) ----------
struct Mesh { uint64_t off; uint32_t first, ntris, nverts, material, model, submesh; };
struct SceneView {
  const uint8_t* buf; clrt_scene_header hdr; std::vector<Mesh> meshes;
  const clrt_light* lights() const { return (const clrt_light*)(buf + sizeof(clrt_scene_header)); }
  const clrt_material* materials() const {
    return (const clrt_material*)(buf + sizeof(clrt_scene_header) + hdr.numberOfLights * 16 + hdr.numberOfSpheres * 16); }
};
SceneView view(const uint8_t* scene) {
  SceneView v; v.buf = scene; memcpy(&v.hdr, scene, sizeof v.hdr);
  size_t off = sizeof(clrt_scene_header) + v.hdr.numberOfLights * 16 + v.hdr.numberOfSpheres * 16 + v.hdr.numberOfMaterials * sizeof(clrt_material);
  uint32_t tri = 0;
  for (uint64_t m = 0; m < v.hdr.numberOfModels; m++) {
    clrt_model_header mh; memcpy(&mh, scene + off, sizeof mh);
    size_t moff = off + sizeof mh;
    for (uint64_t s = 0; s < mh.numberOfSubmeshes; s++) {
      clrt_mesh_header sh; memcpy(&sh, scene + moff, sizeof sh);
      Mesh e = {moff, tri, (uint32_t)sh.numberOfTriangles, (uint32_t)sh.numberOfVertices, (uint32_t)sh.materialIndex, (uint32_t)m, (uint32_t)s};
      if (e.ntris) v.meshes.push_back(e);
      tri += (uint32_t)sh.numberOfTriangles; moff += sh.dataSize;
    }
    off += mh.dataSize;
  }
  return v;
}
inline float ldf(const uint8_t* p) { float f; memcpy(&f, p, 4); return f; }

} // namespace

extern "C" {

// Flatten: triangle i -> {v0,mat},{v1,submesh},{v2,model} + index inside its sub-mesh.
// Equivalent of getTriangleRefByIndex + getVertexAt/getIndexAt (SceneBufferParser.h:197-293).
void oracle_flatten(const uint8_t* scene, clrt_triangle* tris, uint32_t* tri_in_mesh) {
  SceneView sv = view(scene);
  for (size_t k = 0; k < sv.meshes.size(); k++) {
    const Mesh& m = sv.meshes[k];
    const uint8_t* verts = scene + m.off + sizeof(clrt_mesh_header);
    const uint8_t* idx = verts + (size_t)m.nverts * 16;
    for (uint32_t t = 0; t < m.ntris; t++) {
      clrt_triangle& o = tris[m.first + t];
      clrt_float4* dst[3] = {&o.v0, &o.v1, &o.v2};
      const uint32_t ids[3] = {m.material, m.submesh, m.model};
      for (int c = 0; c < 3; c++) {
        uint16_t vi; memcpy(&vi, idx + 2 * (3 * (size_t)t + c), 2);
        const uint8_t* p = verts + (size_t)vi * 16;
        dst[c]->x = ldf(p); dst[c]->y = ldf(p + 4); dst[c]->z = ldf(p + 8);
        memcpy(&dst[c]->w, &ids[c], 4);
      }
      if (tri_in_mesh) tri_in_mesh[m.first + t] = t;
    }
  }
}

// ---- LBVH --------------------------------------------------------------------------------------------------
static inline uint32_t expand_bits(uint32_t v) {           // BVH.h:49-56
  v = (v * 0x00010001u) & 0xFF0000FFu; v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u; v = (v * 0x00000005u) & 0x49249249u; return v;
}
static inline uint32_t morton3d(float x, float y, float z) {   // BVH.h:64-73
  x = rmin(rmax(x * 1024.0f, 0.0f), 1023.0f); y = rmin(rmax(y * 1024.0f, 0.0f), 1023.0f); z = rmin(rmax(z * 1024.0f, 0.0f), 1023.0f);
  return expand_bits((uint32_t)x) * 4 + expand_bits((uint32_t)y) * 2 + expand_bits((uint32_t)z);
}
static inline float normalize_scale(float oldMin, float oldMax, float value) {   // RTKernelUtils.h:63-81 (Q6: no min subtraction)
  const float oldLen = oldMax - oldMin; const float percentage = value / oldLen; const float newLen = 1.0f - 0.0f;
  return percentage * newLen + 0.0f;
}
static inline clrt_aabb triangle_aabb(V3 a, V3 b, V3 c) {   // calculateTriangleAABB, AABB.h:61-79
  clrt_aabb r;
  r.bounds[0].x = rmin(rmin(a.x, b.x), c.x); r.bounds[0].y = rmin(rmin(a.y, b.y), c.y); r.bounds[0].z = rmin(rmin(a.z, b.z), c.z); r.bounds[0].w = 0.0f;
  r.bounds[1].x = rmax(rmax(a.x, b.x), c.x); r.bounds[1].y = rmax(rmax(a.y, b.y), c.y); r.bounds[1].z = rmax(rmax(a.z, b.z), c.z); r.bounds[1].w = 0.0f;
  const uint32_t xFlat = (r.bounds[1].x - r.bounds[0].x) < FLT_EPSILON, yFlat = (r.bounds[1].y - r.bounds[0].y) < FLT_EPSILON, zFlat = (r.bounds[1].z - r.bounds[0].z) < FLT_EPSILON;
  r.bounds[1].x += xFlat * FLT_EPSILON; r.bounds[1].y += yFlat * FLT_EPSILON; r.bounds[1].z += zFlat * FLT_EPSILON;
  r.bounds[0].x -= xFlat * FLT_EPSILON; r.bounds[0].y -= yFlat * FLT_EPSILON; r.bounds[0].z -= zFlat * FLT_EPSILON;
  return r;
}

// calculateMortonCodes (BVHKernels.cl:46-50 -> calculateMorton BVH.h:83-130).  nodes: (2N-1) x 48 B; pairs: N x 8 B.
// Leaf data[0] (parent) is not written by the reference (Q2); here it is set to UINT_MAX like the CUDA path.
void oracle_bvh_morton(const uint8_t* scene, const clrt_triangle* tris, const uint32_t* tri_in_mesh, uint32_t n,
                       clrt_bvh_node* nodes, clrt_uint2* pairs) {
  clrt_scene_header hdr; memcpy(&hdr, scene, sizeof hdr);
  const clrt_aabb bb = hdr.modelsBoundingBox;
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < (long)n; i++) {
    const V3 a = xyz(tris[i].v0), b = xyz(tris[i].v1), c = xyz(tris[i].v2);
    clrt_bvh_node nd;
    nd.boundingBox = triangle_aabb(a, b, c);
    const float cx = (a.x + b.x + c.x) / 3.0f, cy = (a.y + b.y + c.y) / 3.0f, cz = (a.z + b.z + c.z) / 3.0f;
    const float nx = normalize_scale(bb.bounds[0].x, bb.bounds[1].x, cx), ny = normalize_scale(bb.bounds[0].y, bb.bounds[1].y, cy),
                nz = normalize_scale(bb.bounds[0].z, bb.bounds[1].z, cz);
    nd.boundingBox.bounds[0].w = 1.0f;                       // type = LEAF_NODE (BVHData.h:65)
    nd.data[0] = UINT_MAX; nd.data[1] = tri_in_mesh[i];
    memcpy(&nd.data[2], &tris[i].v1.w, 4); memcpy(&nd.data[3], &tris[i].v2.w, 4);
    nodes[i] = nd;
    pairs[i].x = morton3d(nx, ny, nz); pairs[i].y = (uint32_t)i;
  }
}

// Canonical sort: ascending key, equal keys keep input order (stable).
void oracle_sort_pairs_stable(clrt_uint2* pairs, uint64_t n) {
  std::stable_sort(pairs, pairs + n, [](const clrt_uint2& a, const clrt_uint2& b) { return a.x < b.x; });
}
// The reference's bitonic network, literally (BitonicSort.cpp:118-189 host loop; ParallelBitonic_B2 SortKernels.cl:269-287;
// B4/B8 fuse 2/3 of the same stages without changing the result).  n must be 2^k.
int oracle_sort_pairs_bitonic(clrt_uint2* data, uint64_t n) {
  if (n & (n - 1)) return 1;
  for (uint64_t length = 1; length < n; length <<= 1)
    for (uint64_t inc = length; inc > 0; inc >>= 1) {
      const uint64_t dir = length << 1;
      #pragma omp parallel for schedule(static)
      for (long t = 0; t < (long)(n / 2); t++) {
        const uint64_t low = (uint64_t)t & (inc - 1), i = ((uint64_t)t << 1) - low;
        const bool reverse = ((dir & i) == 0);
        const clrt_uint2 x0 = data[i], x1 = data[i + inc];
        const bool swap = reverse ^ (x0.x < x1.x);            // ORDER macro, SortKernels.cl:266
        if (swap) { data[i] = x1; data[i + inc] = x0; }
      }
    }
  return 0;
}
// PrefixSum::computePrefixSum semantics: inclusive (PrefixSumKernels.cl:29-106; Testing/PrefixSumTest.h:66-78 recurrence).
void oracle_scan_inclusive(const uint32_t* in, uint32_t* out, uint64_t n) {
  uint32_t acc = 0;
  for (uint64_t i = 0; i < n; i++) { acc += in[i]; out[i] = acc; }
}

static int find_split(const clrt_uint2* list, int first, int last) {     // BVH.h:138-182
  const uint32_t firstCode = list[first].x, lastCode = list[last].x;
  if (firstCode == lastCode) return first;
  const int commonPrefix = (int)clz32(firstCode ^ lastCode);
  int split = first, step = last - first;
  do {
    step = (step + 1) >> 1;
    const int newSplit = split + step;
    if (newSplit < last) { const int splitPrefix = (int)clz32(firstCode ^ list[newSplit].x); if (splitPrefix > commonPrefix) split = newSplit; }
  } while (step > 1);
  return split;
}
static void determine_range(const clrt_uint2* list, int index, uint32_t size, int* rx, int* ry) {   // BVH.h:190-282
  const int lso = (int)size - 1;
  if (index == 0) { *rx = 0; *ry = lso; return; }
  const int initialindex = index;
  const uint32_t minone = list[index - 1].x, precis = list[index].x, pluone = list[index + 1].x;
  int dir, d_min;
  if (minone == precis && pluone == precis) {
    while (index > 0 && index < lso) { index += 1; if (index >= lso || list[index].x != list[index + 1].x) break; }
    *rx = initialindex; *ry = index; return;
  }
  const int lx = (int)clz32(precis ^ minone), ly = (int)clz32(precis ^ pluone);
  if (lx > ly) { dir = -1; d_min = ly; } else { dir = 1; d_min = lx; }
  int l_max = 2, testindex = index + l_max * dir;
  while ((testindex <= lso && testindex >= 0) ? ((int)clz32(precis ^ list[testindex].x) > d_min) : false) { l_max <<= 1; testindex = index + l_max * dir; }
  int l = 0;
  for (int div = 2; l_max / div >= 1; div <<= 1) {
    const int t = l_max / div, newTest = index + (l + t) * dir;
    if (newTest <= lso && newTest >= 0) { if ((int)clz32(precis ^ list[newTest].x) > d_min) l = l + t; }
  }
  const int other = index + l * dir;
  *rx = std::min(index, other); *ry = std::max(index, other);
}
// buildRadixTree (BVHKernels.cl:55-60 -> constructNode BVH.h:291-321).  literal != 0 keeps the reference's
// `idx < leafCount-2` guard (Q1); 0 builds all N-1 inner nodes (FIXED, what the CUDA path does).
// Inner data[3] and bounds[1].w are never written by the reference; they are zeroed here.
void oracle_bvh_radix_tree(clrt_bvh_node* nodes, const clrt_uint2* list, uint32_t n, int literal) {
  const long cnt = literal ? (long)n - 2 : (long)n - 1;
  #pragma omp parallel for schedule(static)
  for (long idx = 0; idx < cnt; idx++) {
    int rx, ry; determine_range(list, (int)idx, n, &rx, &ry);
    const int split = find_split(list, rx, ry);
    const uint32_t self = (uint32_t)idx + n;
    const bool aIsLeaf = rx == split, bIsLeaf = ry == split + 1;
    const uint32_t cA = aIsLeaf ? list[split].y : (uint32_t)split + n, cB = bIsLeaf ? list[split + 1].y : (uint32_t)split + 1 + n;
    clrt_bvh_node& nd = nodes[self];
    nd.boundingBox.bounds[0].x = nd.boundingBox.bounds[0].y = nd.boundingBox.bounds[0].z = FLT_MAX;
    nd.boundingBox.bounds[1].x = nd.boundingBox.bounds[1].y = nd.boundingBox.bounds[1].z = FLT_MIN;   // sic (Q4)
    nd.boundingBox.bounds[0].w = 0.0f;                       // type = INNER_NODE
    nd.boundingBox.bounds[1].w = 0.0f;
    if (self == n) nd.data[0] = UINT_MAX;
    nd.data[1] = cA; nd.data[2] = cB; nd.data[3] = 0;
    nodes[cA].data[0] = self; nodes[cB].data[0] = self;
  }
}
// computeBoundingBoxes (BVHKernels.cl:65-79) with mergeBoundingBox (BVH.h:328-334) / merge3 (AABB.h:107-113, 4-wide).
// Work-items run one after another; the final boxes are order independent (exact min/max).
void oracle_bvh_refit(clrt_bvh_node* nodes, uint32_t n) {
  std::vector<uint32_t> counters(n ? n : 1, 0u);
  for (uint32_t gid = 0; gid < n; gid++) {
    uint32_t cur = nodes[gid].data[0];
    while (cur != UINT_MAX) {
      if (++counters[cur - n] > 1) {
        const clrt_aabb &a = nodes[nodes[cur].data[1]].boundingBox, &b = nodes[nodes[cur].data[2]].boundingBox, c = nodes[cur].boundingBox;
        clrt_aabb r;
        const float* a0 = &a.bounds[0].x; const float* b0 = &b.bounds[0].x; const float* c0 = &c.bounds[0].x;
        float* r0 = &r.bounds[0].x;
        for (int k = 0; k < 4; k++) r0[k] = rmin(rmin(a0[k], b0[k]), c0[k]);
        for (int k = 4; k < 8; k++) r0[k] = rmax(rmax(a0[k], b0[k]), c0[k]);
        nodes[cur].boundingBox = r;
        cur = nodes[cur].data[0];
      } else break;
    }
  }
}

// ---- rays / traversal ------------------------------------------------------------------------------------------
static inline void gen_ray(const clrt_camera& cam, uint32_t pixelIndex, V3& o, V3& d) {    // CLStructs.h:95-106, Transform.h:282-295
  const float vx = -((float)(pixelIndex % cam.resX) - (float)cam.resX * 0.5f);
  const float vy = (float)(pixelIndex / cam.resX) - (float)cam.resY * 0.5f;
  const float vz = cam.FOVDistance;
  const float* m = cam.viewTransform.data;
  d = normalize(v3(vx * m[0] + vy * m[1] + vz * m[2] + m[3], vx * m[4] + vy * m[5] + vz * m[6] + m[7], vx * m[8] + vy * m[9] + vz * m[10] + m[11]));
  o = v3(m[3], m[7], m[11]);
}
static inline float aabb_intersect(const clrt_aabb& bx, V3 ro, V3 rd) {      // AABB.h:126-142
  const float ix = 1.0f / rd.x, iy = 1.0f / rd.y, iz = 1.0f / rd.z;
  float tmin = (bx.bounds[0.0f > ix].x - ro.x) * ix, tmax = (bx.bounds[1 - (0.0f > ix)].x - ro.x) * ix;
  const float tymin = (bx.bounds[0.0f > iy].y - ro.y) * iy, tymax = (bx.bounds[1 - (0.0f > iy)].y - ro.y) * iy;
  bool flag = !((tmin > tymax) || (tymin > tmax));
  tmin = rmax(tmin, tymin); tmax = rmin(tmax, tymax);
  const float tzmin = (bx.bounds[0.0f > iz].z - ro.z) * iz, tzmax = (bx.bounds[1 - (0.0f > iz)].z - ro.z) * iz;
  flag &= !((tmin > tzmax) || (tzmin > tmax));
  tmin = rmax(tmin, tzmin);
  return tmin * flag;
}
static inline bool point_inside(const clrt_aabb& b, V3 p) {                  // AABB.h:178-184
  return (b.bounds[0].x <= p.x) && (p.x <= b.bounds[1].x) && (b.bounds[0].y <= p.y) && (p.y <= b.bounds[1].y) && (b.bounds[0].z <= p.z) && (p.z <= b.bounds[1].z);
}
static inline float tri_intersect(V3 v0, V3 v1, V3 v2, V3 orig, V3 dir) {    // Triangle.h:66-107 (t*flag; normal separately)
  const V3 e1 = v1 - v0, e2 = v2 - v0, pvec = cross(dir, e2);
  const float det = dot(e1, pvec);
  int flag = fabsf(det) >= FLT_EPSILON;
  const float inv_det = 1.0f / det;
  const V3 tvec = orig - v0;
  const float u = dot(tvec, pvec) * inv_det;
  flag &= (u < 1.0f); flag &= (u >= 0.0f);
  const V3 qvec = cross(tvec, e1);
  const float v = dot(dir, qvec) * inv_det;
  flag &= ((u + v) <= 1.0f); flag &= (v >= 0.0f);
  return dot(e2, qvec) * inv_det * flag;
}
struct TraceStats { uint64_t inner_visits, tri_tests; uint32_t max_stack; };

// bvh_generate_contact (BVH.h:344-408) over the reference node layout: A before B, no pruning, strict `<`.
// pad[0] = primitive id (Q11, additive); miss: everything zero except pad[0] = UINT_MAX.
static clrt_contact trace_one(const clrt_bvh_node* bvh, uint32_t n, const clrt_triangle* tris, V3 o, V3 d, TraceStats* st) {
  std::vector<uint32_t> stack; stack.reserve(64);
  float best = FLT_MAX; uint32_t bestTri = UINT_MAX;
  auto test = [&](uint32_t id) {
    const float t = tri_intersect(xyz(tris[id].v0), xyz(tris[id].v1), xyz(tris[id].v2), o, d);
    if (st) st->tri_tests++;
    if (t > 0 && t < best) { best = t; bestTri = id; }
  };
  if (n == 1) test(0);
  else if (n > 1) {
    uint32_t cur = n;
    stack.push_back(UINT_MAX);
    do {
      const clrt_bvh_node& node = bvh[cur];
      if (node.boundingBox.bounds[0].w == 0.0f) {            // INNER_NODE
        if (st) st->inner_visits++;
        const uint32_t cA = node.data[1], cB = node.data[2];
        const clrt_aabb &bA = bvh[cA].boundingBox, &bB = bvh[cB].boundingBox;
        const float tA = aabb_intersect(bA, o, d), tB = aabb_intersect(bB, o, d);
        const bool aValid = tA > 0 || point_inside(bA, o), bValid = tB > 0 || point_inside(bB, o);
        if (aValid && bValid) { cur = cA; stack.push_back(cB); if (st && stack.size() > st->max_stack) st->max_stack = (uint32_t)stack.size(); }
        else if (aValid) cur = cA;
        else if (bValid) cur = cB;
        else { cur = stack.back(); stack.pop_back(); }
      } else {
        test(cur);                                            // leaf slot id == global triangle index
        cur = stack.back(); stack.pop_back();
      }
    } while (cur != UINT_MAX);
  }
  clrt_contact c; memset(&c, 0, sizeof c);
  c.pad[0] = UINT_MAX;
  if (bestTri != UINT_MAX) {
    const V3 nrm = normalize(cross(xyz(tris[bestTri].v1) - xyz(tris[bestTri].v0), xyz(tris[bestTri].v2) - xyz(tris[bestTri].v0)));
    c.normalAndintersectionDistance.x = nrm.x; c.normalAndintersectionDistance.y = nrm.y; c.normalAndintersectionDistance.z = nrm.z;
    c.normalAndintersectionDistance.w = best;
    memcpy(&c.materialIndex, &tris[bestTri].v0.w, 4);
    c.pad[0] = bestTri;
  }
  return c;
}

// generateContacts (BVHKernels.cl:84-98): contact[pixel] for every pixel.  stats (may be NULL): {inner visits, triangle tests, max stack}.
void oracle_bvh_trace_primary(const clrt_camera* cam, const clrt_bvh_node* bvh, uint32_t n, const clrt_triangle* tris,
                              clrt_contact* out, uint64_t* stats) {
  const long rays = (long)cam->resX * cam->resY;
  uint64_t iv = 0, tt = 0; uint32_t ms = 0;
  #pragma omp parallel for schedule(dynamic, 256) reduction(+:iv, tt) reduction(max:ms)
  for (long i = 0; i < rays; i++) {
    V3 o, d; gen_ray(*cam, (uint32_t)i, o, d);
    TraceStats st = {0, 0, 0};
    clrt_contact c = trace_one(bvh, n, tris, o, d, stats ? &st : nullptr);
    c.pixelIndex = (uint32_t)i;
    out[i] = c;
    iv += st.inner_visits; tt += st.tri_tests; ms = std::max(ms, st.max_stack);
  }
  if (stats) { stats[0] = iv; stats[1] = tt; stats[2] = ms; }
}
// generateContacts2 (BVHKernels.cl:103-117): contact[i] for ray i.
void oracle_bvh_trace_rays(const clrt_ray* rays, uint32_t count, const clrt_bvh_node* bvh, uint32_t n, const clrt_triangle* tris, clrt_contact* out) {
  #pragma omp parallel for schedule(dynamic, 256)
  for (long i = 0; i < (long)count; i++) {
    clrt_contact c = trace_one(bvh, n, tris, xyz(rays[i].origin), xyz(rays[i].direction), nullptr);
    c.pixelIndex = (uint32_t)i;
    out[i] = c;
  }
}
void oracle_generate_primary_rays(const clrt_camera* cam, clrt_ray* rays) {
  const long n = (long)cam->resX * cam->resY;
  for (long i = 0; i < n; i++) {
    V3 o, d; gen_ray(*cam, (uint32_t)i, o, d);
    memset(&rays[i], 0, sizeof(clrt_ray));
    rays[i].idx = (uint32_t)i;
    rays[i].origin.x = o.x; rays[i].origin.y = o.y; rays[i].origin.z = o.z;
    rays[i].direction.x = d.x; rays[i].direction.y = d.y; rays[i].direction.z = d.z;
  }
}

// Demo camera (OpenCLRayTracerDemo.cpp:202-217 + RTFrame :275-276): FOV 90 deg, identity orientation,
// position = (centre x, centre y, minZ - halfZ) of the scene box.  General pose: quaternion (r,i,j,k) + position
// through normalizeQuaternion (Transform.h:88-106) and setOrientationAndPos (:397-413).
void oracle_camera_from_pose(uint32_t w, uint32_t h, float fovDeg, const float* quat, const float* pos3, clrt_camera* out) {
  memset(out, 0, sizeof *out);
  out->resX = w; out->resY = h; out->supersampilngFactor = 1;
  const float PI = 3.14159265359f, DEG2RAD = PI / 180.0f;              // CLStructs.h:60-61
  out->FOVDistance = ((float)h * 0.5f) / tanf(fovDeg * 0.5f * DEG2RAD);  // FOVDistFromAngle :70-74
  float r = quat[0], i = quat[1], j = quat[2], k = quat[3];
  float d = r * r + i * i + j * j + k * k;
  if (d < FLT_EPSILON) { i = j = k = 0.0f; r = 1; }
  else { d = 1.0f / sqrtf(d); r *= d; i *= d; j *= d; k *= d; }
  float* m = out->viewTransform.data;
  m[0] = 1 - (2 * j * j + 2 * k * k); m[1] = 2 * i * j + 2 * k * r; m[2] = 2 * i * k - 2 * j * r; m[3] = pos3[0];
  m[4] = 2 * i * j - 2 * k * r; m[5] = 1 - (2 * i * i + 2 * k * k); m[6] = 2 * j * k + 2 * i * r; m[7] = pos3[1];
  m[8] = 2 * i * k + 2 * j * r; m[9] = 2 * j * k - 2 * i * r; m[10] = 1 - (2 * i * i + 2 * j * j); m[11] = pos3[2];
}
void oracle_demo_camera(const uint8_t* scene, uint32_t w, uint32_t h, clrt_camera* out) {
  clrt_scene_header hdr; memcpy(&hdr, scene, sizeof hdr);
  const clrt_aabb& b = hdr.modelsBoundingBox;
  const float hx = (b.bounds[1].x - b.bounds[0].x) * 0.5f, hy = (b.bounds[1].y - b.bounds[0].y) * 0.5f, hz = (b.bounds[1].z - b.bounds[0].z) * 0.5f;
  const float pos[3] = {b.bounds[0].x + hx, b.bounds[0].y + hy, b.bounds[0].z - hz};
  const float q[4] = {1.0f, 0.0f, 0.0f, 0.0f};
  oracle_camera_from_pose(w, h, 90.0f, q, pos, out);
}

// ---- shading ---------------------------------------------------------------------------------------------------
static const float CAMERA_LIGHT_ENERGY = 100000.0f;       // Shading.h:121
static const float SHADOW_EPS_REL = 1.0e-4f;              // ours (the reference has no shadow rays)

// Shadow rays (ours; restated from shade.cu so the CUDA kernel has a CPU checker): ray r = c*(L+1)+l.
void oracle_shadow_rays(const clrt_camera* cam, const uint8_t* scene, const clrt_contact* contacts, uint32_t n, clrt_ray* rays) {
  SceneView sv = view(scene);
  const uint32_t L = (uint32_t)sv.hdr.numberOfLights, L1 = L + 1;
  for (uint64_t r = 0; r < (uint64_t)n * L1; r++) {
    const uint32_t c = (uint32_t)(r / L1), l = (uint32_t)(r % L1);
    memset(&rays[r], 0, sizeof(clrt_ray)); rays[r].idx = UINT_MAX;
    const clrt_contact& ct = contacts[c];
    if (ct.normalAndintersectionDistance.w > 0.0f && ct.pixelIndex != UINT_MAX) {
      V3 o, d; gen_ray(*cam, ct.pixelIndex, o, d);
      const V3 P = o + d * ct.normalAndintersectionDistance.w;
      const V3 lp = l < L ? xyz(sv.lights()[l].posAndEnergy) : o;
      const V3 dir = normalize(lp - P);
      const V3 org = P + dir * (SHADOW_EPS_REL * ct.normalAndintersectionDistance.w);
      rays[r].idx = (uint32_t)r;
      rays[r].origin.x = org.x; rays[r].origin.y = org.y; rays[r].origin.z = org.z;
      rays[r].direction.x = dir.x; rays[r].direction.y = dir.y; rays[r].direction.z = dir.z;
    }
  }
}

// cpuShadeBlinnPhong (Shading.h:74-166) in the demo's loop (OpenCLRayTracerDemo.cpp:298-316): slot i <- contact i.
// shadow (may be NULL): contacts of the shadow rays above; an occluded light adds nothing (ours).
// rgba8 / rgba_f32 may each be NULL.
void oracle_shade(const clrt_camera* cam, const uint8_t* scene, const clrt_contact* contacts, uint32_t n,
                  const clrt_contact* shadow, uint32_t* rgba8, float* rgba_f32) {
  SceneView sv = view(scene);
  const uint32_t L = (uint32_t)sv.hdr.numberOfLights, L1 = L + 1;
  auto to_u8 = [](float c) { c = (c > 0.0f) ? c : 0.0f; c = (c < 1.0f) ? c : 1.0f; return (uint32_t)(c * 255.0f + 0.5f); };
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < (long)n; i++) {
    const clrt_contact& ct = contacts[i];
    float col[4] = {0, 0, 0, 0}; uint32_t packed = 0;
    if (ct.normalAndintersectionDistance.w > 0.0f && ct.pixelIndex != UINT_MAX) {
      V3 amb = v3(0.2f, 0.2f, 0.2f), dif = v3(1, 1, 1), spc = v3(1, 0, 0); float shininess = 0.4f;
      if (ct.materialIndex < sv.hdr.numberOfMaterials) {
        const clrt_material& m = sv.materials()[ct.materialIndex];
        amb = xyz(m.ambient); dif = xyz(m.diffuse); spc = xyz(m.specular); shininess = m.shininess;
      }
      V3 o, d; gen_ray(*cam, ct.pixelIndex, o, d);
      const float t = ct.normalAndintersectionDistance.w;
      const V3 P = o + d * t, N = xyz(ct.normalAndintersectionDistance), V = normalize(o - P);
      float cr = amb.x, cg = amb.y, cb = amb.z, totalEnergy = 0.0f;
      for (uint32_t l = 0; l < L1; l++) {
        const V3 lp = l < L ? xyz(sv.lights()[l].posAndEnergy) : o; const float e = l < L ? sv.lights()[l].posAndEnergy.w : CAMERA_LIGHT_ENERGY;
        totalEnergy += rmax(1.0f - length3(lp - P) / e, 0.0f) * e;
      }
      for (uint32_t l = 0; l < L1; l++) {
        const V3 lp = l < L ? xyz(sv.lights()[l].posAndEnergy) : o; const float e = l < L ? sv.lights()[l].posAndEnergy.w : CAMERA_LIGHT_ENERGY;
        V3 Lv = lp - P; const float dist = length3(Lv);
        if (shadow) { const float s = shadow[(size_t)i * L1 + l].normalAndintersectionDistance.w; if (s > 0.0f && s < dist - SHADOW_EPS_REL * t) continue; }
        const float intensity = rmax(1.0f - dist / e, 0.0f), pct = (intensity * e) / totalEnergy, ls = intensity * pct;
        Lv = normalize(Lv);
        float LdotN = dot(Lv, N); LdotN = LdotN > 0.0f ? LdotN : 0.0f;
        cr += dif.x * LdotN * ls; cg += dif.y * LdotN * ls; cb += dif.z * LdotN * ls;
        const V3 R = normalize(N * (2.0f * LdotN) - Lv);
        const float ns = powf(dot(R, V), shininess);
        cr += rmax(spc.x * ns * ls, 0.0f); cg += rmax(spc.y * ns * ls, 0.0f); cb += rmax(spc.z * ns * ls, 0.0f);
      }
      col[0] = rmin(cr, 1.0f); col[1] = rmin(cg, 1.0f); col[2] = rmin(cb, 1.0f); col[3] = 1.0f;
      packed = to_u8(col[0]) | (to_u8(col[1]) << 8) | (to_u8(col[2]) << 16) | (255u << 24);
    }
    if (rgba8) rgba8[i] = packed;
    if (rgba_f32) memcpy(rgba_f32 + 4 * i, col, 16);
  }
}

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
void oracle_set_num_threads(int n) {
#ifdef _OPENMP
  omp_set_num_threads(n);
#else
  (void)n;
#endif
}

} // extern "C"
