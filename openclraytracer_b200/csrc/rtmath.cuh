// rtmath.cuh — device arithmetic shared by the build, trace and shading kernels.
//
// The library is compiled with -fmad=false and IEEE division / square root (nvcc defaults), and every expression
// below keeps the reference's operation order, so that results are bit-identical to the reference headers compiled
// for the host (the oracle, -ffp-contract=off).  `rmin`/`rmax` are the reference's host macros
// (R/Include/CLData/CLPortability.h:147-148): compare-select, NOT fminf/fmaxf — they differ on NaN.
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#define RT_FLT_EPSILON 1.192092896e-07f
#define RT_FLT_MAX     3.402823466e+38f
#define RT_FLT_MIN     1.175494351e-38f

__device__ __forceinline__ float rmin(float a, float b) { return (a < b) ? a : b; }
__device__ __forceinline__ float rmax(float a, float b) { return (a > b) ? a : b; }

struct Vec3 { float x, y, z; };
__device__ __forceinline__ Vec3 v3(float x, float y, float z) { Vec3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ Vec3 vsub(Vec3 a, Vec3 b) { return v3(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ Vec3 vadd(Vec3 a, Vec3 b) { return v3(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ Vec3 vscale(Vec3 a, float s) { return v3(a.x * s, a.y * s, a.z * s); }
// CLPortability.h:267-280
__device__ __forceinline__ Vec3 vcross(Vec3 a, Vec3 b) { return v3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
__device__ __forceinline__ float vdot(Vec3 a, Vec3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
// cpu_normalize (CLPortability.h:151-159): l = 1/sqrt(x*x+y*y+z*z); v*l  (exact policy)
__device__ __forceinline__ Vec3 vnormalize(Vec3 a) {
  const float l = 1.0f / sqrtf(a.x * a.x + a.y * a.y + a.z * a.z);
  return v3(a.x * l, a.y * l, a.z * l);
}
__device__ __forceinline__ float vlength(Vec3 a) { return sqrtf(a.x * a.x + a.y * a.y + a.z * a.z); }   // LENGTH3 :279

// generateRay (R/Include/CLData/CLStructs.h:95-106) with transformVectorByMatrix_const (Transform.h:282-295):
// dir = normalize(M * (-(px - resX/2), py - resY/2, FOVDistance) + translation)  [sic, Q9]; origin = translation.
__device__ __forceinline__ void generate_ray(const clrt_camera& cam, uint32_t pixelIndex, Vec3& origin, Vec3& dir) {
  const float vx = -((float)(pixelIndex % cam.resX) - (float)cam.resX * 0.5f);
  const float vy = (float)(pixelIndex / cam.resX) - (float)cam.resY * 0.5f;
  const float vz = cam.FOVDistance;
  const float* m = cam.viewTransform.data;
  Vec3 d = v3(vx * m[0] + vy * m[1] + vz * m[2] + m[3],
              vx * m[4] + vy * m[5] + vz * m[6] + m[7],
              vx * m[8] + vy * m[9] + vz * m[10] + m[11]);
  dir = vnormalize(d);
  origin = v3(m[3], m[7], m[11]);
}

// AABBIntersect (R/Include/CLData/Primitives/AABB.h:126-142): slab test, bounds picked by 0 > 1/d, returns tmin*flag.
// `inv` is 1.0f/d per axis (the reference recomputes the same quotient at every use).
__device__ __forceinline__ float aabb_intersect(float x0, float y0, float z0, float x1, float y1, float z1,
                                                Vec3 ro, Vec3 inv) {
  const bool nx = 0.0f > inv.x, ny = 0.0f > inv.y, nz = 0.0f > inv.z;
  float tmin = ((nx ? x1 : x0) - ro.x) * inv.x;
  float tmax = ((nx ? x0 : x1) - ro.x) * inv.x;
  const float tymin = ((ny ? y1 : y0) - ro.y) * inv.y;
  const float tymax = ((ny ? y0 : y1) - ro.y) * inv.y;
  bool flag = !((tmin > tymax) || (tymin > tmax));
  tmin = rmax(tmin, tymin);
  tmax = rmin(tmax, tymax);
  const float tzmin = ((nz ? z1 : z0) - ro.z) * inv.z;
  const float tzmax = ((nz ? z0 : z1) - ro.z) * inv.z;
  flag &= !((tmin > tzmax) || (tzmin > tmax));
  tmin = rmax(tmin, tzmin);
  return tmin * (float)flag;
}
// isPointInside (AABB.h:178-184)
__device__ __forceinline__ bool point_inside(float x0, float y0, float z0, float x1, float y1, float z1, Vec3 p) {
  return (x0 <= p.x) && (p.x <= x1) && (y0 <= p.y) && (p.y <= y1) && (z0 <= p.z) && (p.z <= z1);
}

// triangleIntersect (R/Include/CLData/Primitives/Triangle.h:66-107): Moeller-Trumbore, two sided, branch-free flags.
// Returns t*flag exactly as the reference (NaN/0 when rejected); the normal is computed by the caller for the
// winning triangle only (same value: it does not depend on the ray).
__device__ __forceinline__ float triangle_intersect(Vec3 v0, Vec3 v1, Vec3 v2, Vec3 orig, Vec3 dir) {
  const Vec3 e1 = vsub(v1, v0), e2 = vsub(v2, v0);
  const Vec3 pvec = vcross(dir, e2);
  const float det = vdot(e1, pvec);
  int flag = fabsf(det) >= RT_FLT_EPSILON;
  const float inv_det = 1.0f / det;
  const Vec3 tvec = vsub(orig, v0);
  const float u = vdot(tvec, pvec) * inv_det;
  flag &= (u < 1.0f);
  flag &= (u >= 0.0f);
  const Vec3 qvec = vcross(tvec, e1);
  const float v = vdot(dir, qvec) * inv_det;
  flag &= ((u + v) <= 1.0f);
  flag &= (v >= 0.0f);
  return vdot(e2, qvec) * inv_det * (float)flag;
}
__device__ __forceinline__ Vec3 triangle_normal(Vec3 v0, Vec3 v1, Vec3 v2) {
  return vnormalize(vcross(vsub(v1, v0), vsub(v2, v0)));
}

// expandBits / morton3D (BVH.h:49-73)
__device__ __forceinline__ uint32_t expand_bits(uint32_t v) {
  v = (v * 0x00010001u) & 0xFF0000FFu;
  v = (v * 0x00000101u) & 0x0F00F00Fu;
  v = (v * 0x00000011u) & 0xC30C30C3u;
  v = (v * 0x00000005u) & 0x49249249u;
  return v;
}
__device__ __forceinline__ uint32_t morton3d(float x, float y, float z) {
  x = rmin(rmax(x * 1024.0f, 0.0f), 1023.0f);
  y = rmin(rmax(y * 1024.0f, 0.0f), 1023.0f);
  z = rmin(rmax(z * 1024.0f, 0.0f), 1023.0f);
  return expand_bits((uint32_t)x) * 4 + expand_bits((uint32_t)y) * 2 + expand_bits((uint32_t)z);
}
// normalizeScale -> translateScale(oldMin, oldMax, value, 0, 1)  (RTKernelUtils.h:63-81): no min subtraction (Q6)
__device__ __forceinline__ float normalize_scale(float oldMin, float oldMax, float value) {
  const float oldLen = oldMax - oldMin;
  const float percentage = value / oldLen;
  const float newLen = 1.0f - 0.0f;
  return percentage * newLen + 0.0f;
}
// calculateMorton (BVH.h:83-130): centroid (v1+v2+v3)/3.0f normalised to the scene box WITHOUT subtracting its minimum (Q6), 10 bits
// per axis.  Shared by the LBVH build (bvh.cu) and by the tie-break of the structures the reference does not have (kd.cu, the corrected
// grid walk): among hits with bit-equal t the reference's traversal keeps the triangle that comes first in ITS sorted order, i.e. the
// smaller (Morton code, triangle index).
__device__ __forceinline__ uint32_t reference_morton(float4 a, float4 b, float4 c, const clrt_aabb& sceneBB) {
  const float cx = (a.x + b.x + c.x) / 3.0f, cy = (a.y + b.y + c.y) / 3.0f, cz = (a.z + b.z + c.z) / 3.0f;
  const float nx = normalize_scale(sceneBB.bounds[0].x, sceneBB.bounds[1].x, cx);
  const float ny = normalize_scale(sceneBB.bounds[0].y, sceneBB.bounds[1].y, cy);
  const float nz = normalize_scale(sceneBB.bounds[0].z, sceneBB.bounds[1].z, cz);
  return morton3d(nx, ny, nz);
}
// true when triangle `id` precedes triangle `other` in the reference's sorted (Morton code, index) order
__device__ __forceinline__ bool reference_order_before(const clrt_triangle* __restrict__ tris, uint32_t id, uint32_t other, const clrt_aabb& sceneBB) {
  const float4* p = reinterpret_cast<const float4*>(tris + id); const float4* q = reinterpret_cast<const float4*>(tris + other);
  const uint32_t ci = reference_morton(__ldg(p), __ldg(p + 1), __ldg(p + 2), sceneBB), co = reference_morton(__ldg(q), __ldg(q + 1), __ldg(q + 2), sceneBB);
  return ci < co || (ci == co && id < other);
}
