// shade_math.cuh — shadow-ray construction and Blinn-Phong shading shared by the stand-alone kernels (shade.cu) and the
// fused shadow+shade kernel (trace.cu).
//  cpuShadeBlinnPhong     R/Include/CLData/Shading.h:74-166 (host loop R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:298-316)
//  lightEnergyPercentage  R/Include/CLData/Primitives/Light.h:55-58
#pragma once
#include "common.cuh"
#include "rtmath.cuh"

constexpr float CAMERA_LIGHT_ENERGY = 100000.0f;        // Shading.h:121
constexpr float SHADOW_EPS_REL = 1.0e-4f;               // shadow-ray origin offset, relative to the primary hit distance

__device__ __forceinline__ Vec3 light_position(const clrt_light* __restrict__ lights, uint32_t n_lights, uint32_t l, Vec3 camPos, float& energy) {
  if (l < n_lights) { const float4 q = __ldg(reinterpret_cast<const float4*>(lights + l)); energy = q.w; return v3(q.x, q.y, q.z); }
  energy = CAMERA_LIGHT_ENERGY; return camPos;           // the synthetic camera light (Shading.h:114-126)
}

__device__ __forceinline__ uint32_t to_u8(float c) {
  c = (c > 0.0f) ? c : 0.0f;            // NaN -> 0
  c = (c < 1.0f) ? c : 1.0f;
  return (uint32_t)(c * 255.0f + 0.5f);
}

// Shadow ray of light l for the contact (pixel head.x, hit distance t): from the hit point, pushed SHADOW_EPS_REL*t towards the
// light, to the light.  `limit` = the distance below which a hit occludes the light (shade: st < dist - eps*t).
__device__ __forceinline__ void shadow_ray(const clrt_camera& cam, uint32_t pixel, float t, const clrt_light* __restrict__ lights,
                                           uint32_t n_lights, uint32_t l, Vec3& org, Vec3& dir) {
  Vec3 o, d; generate_ray(cam, pixel, o, d);
  const Vec3 P = vadd(o, vscale(d, t));
  float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
  dir = vnormalize(vsub(lp, P));
  org = vadd(P, vscale(dir, SHADOW_EPS_REL * t));
}

// Blinn-Phong of one contact.  occluded(l) tells whether light l is shadowed (ours; the reference has no shadows).
template <typename Occluded>
__device__ __forceinline__ float4 shade_contact(const clrt_camera& cam, uint4 head, float4 nd, const clrt_light* __restrict__ lights,
                                                uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
                                                Occluded occluded, uint32_t& packed) {
  float4 color = make_float4(0.f, 0.f, 0.f, 0.f);
  packed = 0u;
  if (nd.w > 0.0f && head.x != 0xFFFFFFFFu) {                      // OpenCLRayTracerDemo.cpp:305
    // default material (Shading.h:76-86) unless the index is valid (:92-96)
    Vec3 amb = v3(0.2f, 0.2f, 0.2f), dif = v3(1.f, 1.f, 1.f), spc = v3(1.f, 0.f, 0.f);
    float shininess = 0.4f;
    if (head.y < n_materials) {
      const float4* m = reinterpret_cast<const float4*>(materials + head.y);
      const float4 a = __ldg(m), d4 = __ldg(m + 1), s4 = __ldg(m + 2), tail = __ldg(m + 5);
      amb = v3(a.x, a.y, a.z); dif = v3(d4.x, d4.y, d4.z); spc = v3(s4.x, s4.y, s4.z); shininess = tail.x;
    }
    Vec3 o, d; generate_ray(cam, head.x, o, d);
    const Vec3 P = vadd(o, vscale(d, nd.w));
    const Vec3 N = v3(nd.x, nd.y, nd.z);
    const Vec3 V = vnormalize(vsub(o, P));
    float cr = amb.x, cg = amb.y, cb = amb.z;
    const uint32_t L1 = n_lights + 1;
    float totalEnergy = 0.0f;
    for (uint32_t l = 0; l < L1; l++) {
      float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
      const float dist = vlength(vsub(lp, P));
      totalEnergy += rmax(1.0f - dist / energy, 0.0f) * energy;
    }
    for (uint32_t l = 0; l < L1; l++) {
      float energy; const Vec3 lp = light_position(lights, n_lights, l, o, energy);
      Vec3 Lv = vsub(lp, P);
      const float dist = vlength(Lv);
      if (occluded(l, dist)) continue;
      const float intensity = rmax(1.0f - dist / energy, 0.0f);
      const float pct = (intensity * energy) / totalEnergy;
      const float ls = intensity * pct;
      Lv = vnormalize(Lv);
      float LdotN = vdot(Lv, N);
      LdotN = LdotN > 0.0f ? LdotN : 0.0f;
      cr += dif.x * LdotN * ls; cg += dif.y * LdotN * ls; cb += dif.z * LdotN * ls;
      const Vec3 R = vnormalize(vsub(vscale(N, 2.0f * LdotN), Lv));
      const float nSpecular = powf(vdot(R, V), shininess);
      cr += rmax(spc.x * nSpecular * ls, 0.0f); cg += rmax(spc.y * nSpecular * ls, 0.0f); cb += rmax(spc.z * nSpecular * ls, 0.0f);
    }
    color = make_float4(rmin(cr, 1.0f), rmin(cg, 1.0f), rmin(cb, 1.0f), 1.0f);
    packed = to_u8(color.x) | (to_u8(color.y) << 8) | (to_u8(color.z) << 16) | (255u << 24);
  }
  return color;
}
