// shade.cu — ray generation, shadow rays and "simple shading" as CUDA kernels.
//  generateRay            R/Include/CLData/CLStructs.h:95-106
//  cpuShadeBlinnPhong     R/Include/CLData/Shading.h:74-166 (a HOST loop in the reference:
//                         R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:298-316)
//  lightEnergyPercentage  R/Include/CLData/Primitives/Light.h:55-58
// Shadow rays do not exist in the reference (KNOWN_ISSUES_AND_TODO.txt:27-30); they are produced here and traced
// through the reference's generic-ray entry point (generateContacts(rays, contacts, n)).
#include "shade_math.cuh"

namespace {

__global__ void __launch_bounds__(256)
primary_rays_kernel(clrt_camera cam, clrt_ray* __restrict__ rays) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cam.resX * cam.resY) return;
  Vec3 o, d; generate_ray(cam, i, o, d);
  uint4* p = reinterpret_cast<uint4*>(rays + i);
  p[0] = make_uint4(i, 0u, 0u, 0u);
  reinterpret_cast<float4*>(p)[1] = make_float4(o.x, o.y, o.z, 0.0f);
  reinterpret_cast<float4*>(p)[2] = make_float4(d.x, d.y, d.z, 0.0f);
}

// One thread per (contact, light).  Ray index r = c*(L+1)+l.
__global__ void __launch_bounds__(256)
shadow_rays_kernel(clrt_camera cam, const clrt_contact* __restrict__ contacts, uint32_t n, const clrt_light* __restrict__ lights,
                   uint32_t n_lights, clrt_ray* __restrict__ rays, unsigned long long* __restrict__ ray_counter) {
  const uint32_t L1 = n_lights + 1;
  const uint64_t r = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x;
  const bool in_range = r < (uint64_t)n * L1;
  bool live = false;
  if (in_range) {
    const uint32_t c = (uint32_t)(r / L1), l = (uint32_t)(r % L1);
    const uint4 head = __ldg(reinterpret_cast<const uint4*>(contacts + c));
    const float4 nd = __ldg(reinterpret_cast<const float4*>(contacts + c) + 1);
    uint4 h = make_uint4(0xFFFFFFFFu, 0u, 0u, 0u);
    float4 ro = make_float4(0.f, 0.f, 0.f, 0.f), rd = make_float4(0.f, 0.f, 0.f, 0.f);
    if (nd.w > 0.0f && head.x != 0xFFFFFFFFu) {
      Vec3 org, dir; shadow_ray(cam, head.x, nd.w, lights, n_lights, l, org, dir);
      h.x = (uint32_t)r; ro = make_float4(org.x, org.y, org.z, 0.f); rd = make_float4(dir.x, dir.y, dir.z, 0.f);
      live = true;
    }
    uint4* p = reinterpret_cast<uint4*>(rays + r);
    p[0] = h; reinterpret_cast<float4*>(p)[1] = ro; reinterpret_cast<float4*>(p)[2] = rd;
  }
  const unsigned m = __ballot_sync(0xffffffffu, live);
  if ((threadIdx.x & 31) == 0 && m && ray_counter) atomicAdd(ray_counter, (unsigned long long)__popc(m));
}

// Incoherent secondary rays (BASELINE.json configs[3]; the reference's TODO list names them,
// KNOWN_ISSUES_AND_TODO.txt:27-33): one cosine-weighted diffuse bounce per primary hit.  The two uniforms come from
// splitmix64(seed * 0x9E3779B97F4A7C15 + pixelIndex) (high / low 24 bits), the direction is Malley's: a uniform disk point
// lifted to the hemisphere around the geometric normal turned against the incoming ray; the origin is pushed 1e-4 * t along
// that normal (the shadow rays' offset).  Misses give the zero-direction ray the trace kernels answer without traversal.
__device__ __forceinline__ uint64_t splitmix64_dev(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull;
  x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
  x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
  return x ^ (x >> 31);
}
// Secondary rays (the reference's TODO 2.2, KNOWN_ISSUES_AND_TODO.txt: "generating rays from viewing rays-scene contacts, and execute
// ray tracing kernel in loop as many times as specified by ray tracing depth").  Contact i is either a primary contact (incoming
// ray = generateRay(camera, pixelIndex)) or the contact of in_rays[i] from the generic-ray entry point; the new ray carries the pixel
// along in idx.  KIND 0: cosine-weighted diffuse direction about the normal; KIND 1: mirror direction d - 2 (d.n) n.
template <int KIND, bool FROM_RAYS>
__global__ void __launch_bounds__(256)
secondary_rays_kernel(clrt_camera cam, const clrt_ray* in_rays, const clrt_contact* __restrict__ contacts, uint32_t n, uint64_t seed,
                      clrt_ray* out_rays) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 head = __ldg(reinterpret_cast<const uint4*>(contacts + i));
  const float4 nd = __ldg(reinterpret_cast<const float4*>(contacts + i) + 1);
  uint4 h = make_uint4(0xFFFFFFFFu, 0u, 0u, 0u);
  float4 ro = make_float4(0.f, 0.f, 0.f, 0.f), rd = make_float4(0.f, 0.f, 0.f, 0.f);
  Vec3 o = v3(0.f, 0.f, 0.f), d = v3(0.f, 0.f, 0.f);
  uint32_t pixel = 0xFFFFFFFFu;
  if (FROM_RAYS) {                                          // in_rays may alias out_rays: ray i is read before it is written, by this thread only
    const uint4 rh = *reinterpret_cast<const uint4*>(in_rays + i);
    const float4 io = reinterpret_cast<const float4*>(in_rays + i)[1], id = reinterpret_cast<const float4*>(in_rays + i)[2];
    pixel = rh.x; o = v3(io.x, io.y, io.z); d = v3(id.x, id.y, id.z);
  } else if (head.x != 0xFFFFFFFFu) {
    pixel = head.x; generate_ray(cam, pixel, o, d);
  }
  if (nd.w > 0.0f && pixel != 0xFFFFFFFFu) {
    Vec3 nrm = v3(nd.x, nd.y, nd.z);
    const float dn = vdot(nrm, d);
    if (dn > 0.0f) nrm = vscale(nrm, -1.0f);                                   // face the incoming ray (two-sided triangles)
    const Vec3 P = vadd(o, vscale(d, nd.w));
    Vec3 dir;
    if (KIND == 1) {
      dir = vnormalize(vsub(d, vscale(nrm, 2.0f * vdot(nrm, d))));
    } else {
      const uint64_t u = splitmix64_dev(seed * 0x9E3779B97F4A7C15ull + (uint64_t)pixel);
      const float u1 = (float)(uint32_t)(u >> 40) * (1.0f / 16777216.0f), u2 = (float)(uint32_t)(u & 0xFFFFFFu) * (1.0f / 16777216.0f);
      const float r = sqrtf(u1), phi = 6.283185307179586f * u2;
      float sn, cs; sincosf(phi, &sn, &cs);
      const float dx = r * cs, dy = r * sn, dz = sqrtf(fmaxf(0.0f, 1.0f - u1));
      // orthonormal basis around the normal: the axis least aligned with it, crossed twice
      const Vec3 ax = fabsf(nrm.x) < 0.5f ? v3(1.0f, 0.0f, 0.0f) : v3(0.0f, 1.0f, 0.0f);
      const Vec3 t1 = vnormalize(vcross(ax, nrm)), t2 = vcross(nrm, t1);
      dir = vnormalize(vadd(vadd(vscale(t1, dx), vscale(t2, dy)), vscale(nrm, dz)));
    }
    const Vec3 org = vadd(P, vscale(nrm, SHADOW_EPS_REL * nd.w));
    h.x = pixel; ro = make_float4(org.x, org.y, org.z, 0.f); rd = make_float4(dir.x, dir.y, dir.z, 0.f);
  }
  uint4* p = reinterpret_cast<uint4*>(out_rays + i);
  p[0] = h; reinterpret_cast<float4*>(p)[1] = ro; reinterpret_cast<float4*>(p)[2] = rd;
}

// One thread per contact.
__global__ void __launch_bounds__(256)
shade_kernel(clrt_camera cam, const clrt_contact* __restrict__ contacts, uint32_t n, const clrt_light* __restrict__ lights,
             uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
             const clrt_contact* __restrict__ shadow, uint32_t* __restrict__ rgba8, float4* __restrict__ rgba_f32) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + i));
  const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + i) + 1);
  const uint32_t L1 = n_lights + 1;
  uint32_t packed;
  const float4 color = shade_contact(cam, head, nd, lights, n_lights, materials, n_materials,
    [&](uint32_t l, float dist) {                                  // ours: occluded lights add nothing
      if (!shadow) return false;
      const float st = __ldcs(&shadow[(size_t)i * L1 + l].normalAndintersectionDistance.w);
      return st > 0.0f && st < dist - SHADOW_EPS_REL * nd.w;
    }, packed);
  if (rgba8) __stcs(rgba8 + i, packed);
  if (rgba_f32) __stcs(rgba_f32 + i, color);
}

// Bounce frame (BASELINE.json configs[3]: one diffuse bounce per primary hit): Blinn-Phong of the primary contact, halved when the
// bounce ray of the same slot hit something (one-sample ambient occlusion).  Ours; the reference has no secondary rays (its TODO 2.2).
__global__ void __launch_bounds__(256)
shade_bounce_kernel(clrt_camera cam, const clrt_contact* __restrict__ contacts, uint32_t n, const clrt_light* __restrict__ lights,
                    uint32_t n_lights, const clrt_material* __restrict__ materials, uint32_t n_materials,
                    const clrt_contact* __restrict__ bounce, uint32_t* __restrict__ rgba8) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint4 head = __ldcs(reinterpret_cast<const uint4*>(contacts + i));
  const float4 nd = __ldcs(reinterpret_cast<const float4*>(contacts + i) + 1);
  uint32_t packed;
  shade_contact(cam, head, nd, lights, n_lights, materials, n_materials, [](uint32_t, float) { return false; }, packed);
  if (packed && __ldcs(&bounce[i].normalAndintersectionDistance.w) > 0.0f)
    packed = ((packed >> 1) & 0x007F7F7Fu) | (packed & 0xFF000000u);
  __stcs(rgba8 + i, packed);
}

// stripes (rank-major, each padded to ceil(tiles/nranks) tiles, tile-major, 256 px per tile) -> raster
__global__ void __launch_bounds__(256)
untile_kernel(const uint32_t* __restrict__ stripes, uint32_t resX, uint32_t resY, uint32_t nranks, uint32_t* __restrict__ raster) {
  const TileGrid g = make_tile_grid(resX, resY);
  const uint32_t gt = blockIdx.x;                         // global tile
  const uint32_t rank = gt % nranks, lt = gt / nranks;
  // every rank's stripe is padded to ceil(tiles/nranks) tiles so that an all-gather of equal chunks lays them out
  const uint32_t stripe_first_tile = rank * ((g.tiles + nranks - 1) / nranks);
  const uint32_t t = threadIdx.x;
  const uint32_t x = (gt % g.tilesX) * CLRT_TILE_W + (t & 31u), y = (gt / g.tilesX) * CLRT_TILE_H + (t >> 5);
  if (x < resX && y < resY) raster[y * resX + x] = __ldcs(stripes + (size_t)(stripe_first_tile + lt) * CLRT_TILE_PIXELS + t);
}

// s x s box filter over an (resX*s) x (resY*s) RGBA8 raster -> resX x resY; per channel (sum + s*s/2) / (s*s)
__global__ void __launch_bounds__(256)
downsample_kernel(const uint32_t* __restrict__ hi, uint32_t resX, uint32_t resY, uint32_t s, uint32_t* __restrict__ lo) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= resX * resY) return;
  const uint32_t x = i % resX, y = i / resX;
  uint32_t r = 0, g = 0, b = 0, a = 0;
  for (uint32_t j = 0; j < s; j++)
    for (uint32_t k = 0; k < s; k++) {
      const uint32_t v = __ldcs(hi + (size_t)(y * s + j) * (resX * s) + (x * s + k));
      r += v & 255u; g += (v >> 8) & 255u; b += (v >> 16) & 255u; a += v >> 24;
    }
  const uint32_t q = s * s, h = q / 2;
  lo[i] = ((r + h) / q) | (((g + h) / q) << 8) | (((b + h) / q) << 16) | (((a + h) / q) << 24);
}

} // namespace

int shade_launch_downsample(clrt_ctx* ctx, const uint32_t* hi, uint32_t resX, uint32_t resY, uint32_t s, uint32_t* lo) {
  const uint32_t n = resX * resY;
  if (!n) return 0;
  downsample_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(hi, resX, resY, s, lo);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_generate_primary_rays(clrt_ctx* ctx, const clrt_camera* h_camera, void* dptr_rays) {
  if (!ctx || !h_camera || !dptr_rays) CLRT_FAIL("clrt_generate_primary_rays: NULL argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  const uint64_t n = (uint64_t)h_camera->resX * h_camera->resY;
  if (n == 0) return 0;
  if (n > 0xFFFFFFFFull) CLRT_FAIL("clrt_generate_primary_rays: resolution too large");
  primary_rays_kernel<<<(unsigned)((n + 255) / 256), 256, 0, ctx->stream>>>(*h_camera, (clrt_ray*)dptr_rays);
  CLRT_CHECK_LAUNCH();
  return 0;
}

int shade_launch_shadow_rays(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera& cam, const clrt_contact* contacts,
                             uint32_t n, clrt_ray* rays, unsigned long long* d_counter) {
  const uint64_t total = (uint64_t)n * (scene->n_lights + 1);
  if (!total) return 0;
  if (total > 0xFFFFFFFFull) CLRT_FAIL("shadow rays: %llu rays exceed 2^32-1", (unsigned long long)total);
  shadow_rays_kernel<<<(unsigned)((total + 255) / 256), 256, 0, ctx->stream>>>(cam, contacts, n, scene->d_lights, scene->n_lights, rays, d_counter);
  CLRT_CHECK_LAUNCH();
  return 0;
}

int shade_launch(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera& cam, const clrt_contact* contacts, uint32_t n,
                 const clrt_contact* shadow, uint32_t* rgba8, float4* rgba_f32, cudaStream_t stream) {
  if (!n) return 0;
  shade_kernel<<<(n + 255) / 256, 256, 0, stream ? stream : ctx->stream>>>(cam, contacts, n, scene->d_lights, scene->n_lights, scene->d_materials,
                                                        scene->n_materials, shadow, rgba8, rgba_f32);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_generate_shadow_rays(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera,
                                         const void* dptr_contacts, uint32_t n, void* dptr_rays, uint32_t* out_rays_per_contact) {
  if (!ctx || !scene || !h_camera || (n && (!dptr_contacts || !dptr_rays))) CLRT_FAIL("clrt_generate_shadow_rays: NULL argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (out_rays_per_contact) *out_rays_per_contact = scene->n_lights + 1;
  return shade_launch_shadow_rays(ctx, scene, *h_camera, (const clrt_contact*)dptr_contacts, n, (clrt_ray*)dptr_rays, nullptr);
}

extern "C" int clrt_generate_secondary_rays(clrt_ctx* ctx, int kind, const clrt_camera* h_camera, const void* dptr_in_rays,
                                            const void* dptr_contacts, uint32_t n, uint64_t seed, void* dptr_rays) {
  if (!ctx || (!h_camera && !dptr_in_rays) || (n && (!dptr_contacts || !dptr_rays))) CLRT_FAIL("clrt_generate_secondary_rays: NULL argument");
  if (kind != CLRT_RAYS_DIFFUSE && kind != CLRT_RAYS_REFLECT) CLRT_FAIL("clrt_generate_secondary_rays: kind %d", kind);
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (!n) return 0;
  clrt_camera cam; memset(&cam, 0, sizeof cam);
  if (h_camera) cam = *h_camera;
  const clrt_ray* in = (const clrt_ray*)dptr_in_rays; const clrt_contact* c = (const clrt_contact*)dptr_contacts; clrt_ray* out = (clrt_ray*)dptr_rays;
  const unsigned grid = (n + 255) / 256;
  if (kind == CLRT_RAYS_DIFFUSE) {
    if (in) secondary_rays_kernel<0, true><<<grid, 256, 0, ctx->stream>>>(cam, in, c, n, seed, out);
    else    secondary_rays_kernel<0, false><<<grid, 256, 0, ctx->stream>>>(cam, in, c, n, seed, out);
  } else {
    if (in) secondary_rays_kernel<1, true><<<grid, 256, 0, ctx->stream>>>(cam, in, c, n, seed, out);
    else    secondary_rays_kernel<1, false><<<grid, 256, 0, ctx->stream>>>(cam, in, c, n, seed, out);
  }
  CLRT_CHECK_LAUNCH();
  return 0;
}
extern "C" int clrt_generate_bounce_rays(clrt_ctx* ctx, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n, uint64_t seed,
                                         void* dptr_rays) {
  if (!h_camera) CLRT_FAIL("clrt_generate_bounce_rays: NULL argument");
  return clrt_generate_secondary_rays(ctx, CLRT_RAYS_DIFFUSE, h_camera, nullptr, dptr_contacts, n, seed, dptr_rays);
}

extern "C" int clrt_shade(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera, const void* dptr_contacts,
                          uint32_t n, const void* dptr_shadow_contacts, void* dptr_rgba8, void* dptr_rgba_f32) {
  if (!ctx || !scene || !h_camera || (n && !dptr_contacts)) CLRT_FAIL("clrt_shade: NULL argument");
  if (!dptr_rgba8 && !dptr_rgba_f32) CLRT_FAIL("clrt_shade: no output buffer");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  return shade_launch(ctx, scene, *h_camera, (const clrt_contact*)dptr_contacts, n, (const clrt_contact*)dptr_shadow_contacts,
                      (uint32_t*)dptr_rgba8, (float4*)dptr_rgba_f32, nullptr);
}

extern "C" int clrt_shade_bounce(clrt_ctx* ctx, const clrt_scene* scene, const clrt_camera* h_camera, const void* dptr_contacts, uint32_t n,
                                 const void* dptr_bounce_contacts, void* dptr_rgba8) {
  if (!ctx || !scene || !h_camera || (n && (!dptr_contacts || !dptr_bounce_contacts || !dptr_rgba8))) CLRT_FAIL("clrt_shade_bounce: NULL argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  if (!n) return 0;
  shade_bounce_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(*h_camera, (const clrt_contact*)dptr_contacts, n, scene->d_lights, scene->n_lights,
      scene->d_materials, scene->n_materials, (const clrt_contact*)dptr_bounce_contacts, (uint32_t*)dptr_rgba8);
  CLRT_CHECK_LAUNCH();
  return 0;
}

extern "C" int clrt_frame_untile_on(clrt_ctx* ctx, void* stream, const void* dptr_stripes, uint32_t resX, uint32_t resY, uint32_t nranks,
                                    void* dptr_raster_rgba8) {
  if (!ctx || !dptr_stripes || !dptr_raster_rgba8 || !nranks) CLRT_FAIL("clrt_frame_untile: bad argument");
  CLRT_CUDA(cudaSetDevice(ctx->device));
  const TileGrid g = make_tile_grid(resX, resY);
  if (!g.tiles) return 0;
  untile_kernel<<<g.tiles, 256, 0, (cudaStream_t)stream>>>((const uint32_t*)dptr_stripes, resX, resY, nranks, (uint32_t*)dptr_raster_rgba8);
  CLRT_CHECK_LAUNCH();
  return 0;
}
extern "C" int clrt_frame_untile(clrt_ctx* ctx, const void* dptr_stripes, uint32_t resX, uint32_t resY, uint32_t nranks, void* dptr_raster_rgba8) {
  if (!ctx) CLRT_FAIL("clrt_frame_untile: bad argument");
  return clrt_frame_untile_on(ctx, ctx->stream, dptr_stripes, resX, resY, nranks, dptr_raster_rgba8);
}
