// scene.cu — Scene::loadToGPU (R/src/CLRayTracerEngine/Scene.cpp:451-456) + device-side flattening.
//
// The packed v0.6 scene buffer (R/Include/CLData/SceneBufferParser.h: SceneHeader | lights | spheres | materials |
// models -> sub-meshes {MeshHeader, float3[nv] (16 B each), ushort[ni]}) is uploaded unchanged.  The reference's
// kernels re-walk the model/sub-mesh headers for EVERY triangle fetch (getTriangleRefByIndex :253-293,
// getModelAtIndex/getMeshAtIndex :175-195) and read float3 arrays that are not 16-byte aligned (mesh blocks are
// not padded: Scene.cpp:80-86).  Here the header walk happens once on the host at upload (a table of
// sub-meshes), and one kernel gathers every triangle into a 16-byte aligned clrt_triangle[N] (48 B) that all later
// kernels read.  The gather reads the source with 2-byte loads because mesh blocks are only 2-byte aligned.
#include "common.cuh"

namespace {

__device__ __forceinline__ float load_f32_2b(const uint8_t* p) {      // p is 2-byte aligned
  const uint16_t lo = *reinterpret_cast<const uint16_t*>(p), hi = *reinterpret_cast<const uint16_t*>(p + 2);
  return __uint_as_float((uint32_t)lo | ((uint32_t)hi << 16));
}

__global__ void __launch_bounds__(256)
flatten_triangles_kernel(const uint8_t* __restrict__ scene, const MeshEntry* __restrict__ meshes, uint32_t n_meshes,
                         uint32_t n_tris, clrt_triangle* __restrict__ tris, uint32_t* __restrict__ tri_in_mesh) {
  const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_tris) return;
  // binary search: last mesh with firstTriangle <= i  (empty sub-meshes share a firstTriangle; pick the one that owns i)
  uint32_t lo = 0, hi = n_meshes;
  while (hi - lo > 1) { const uint32_t mid = (lo + hi) >> 1; if (meshes[mid].firstTriangle <= i) lo = mid; else hi = mid; }
  const MeshEntry m = meshes[lo];
  const uint32_t t = i - m.firstTriangle;
  const uint8_t* mesh = scene + m.byteOffset;
  const uint8_t* verts = mesh + sizeof(clrt_mesh_header);
  const uint16_t* idx = reinterpret_cast<const uint16_t*>(verts + (size_t)m.numVertices * 16);
  float v[3][3];
#pragma unroll
  for (int k = 0; k < 3; k++) {
    const uint8_t* p = verts + (size_t)idx[3 * t + k] * 16;
    v[k][0] = load_f32_2b(p); v[k][1] = load_f32_2b(p + 4); v[k][2] = load_f32_2b(p + 8);
  }
  float4* o = reinterpret_cast<float4*>(tris + i);
  o[0] = make_float4(v[0][0], v[0][1], v[0][2], __uint_as_float(m.materialIndex));
  o[1] = make_float4(v[1][0], v[1][1], v[1][2], __uint_as_float(m.submeshIndex));
  o[2] = make_float4(v[2][0], v[2][1], v[2][2], __uint_as_float(m.modelIndex));
  tri_in_mesh[i] = t;
}

} // namespace

extern "C" int clrt_scene_upload(clrt_ctx* ctx, const void* h_scene, size_t bytes, clrt_scene** out) {
  if (!ctx || !h_scene || !out) CLRT_FAIL("clrt_scene_upload: NULL argument");
  if (bytes < sizeof(clrt_scene_header)) CLRT_FAIL("clrt_scene_upload: %zu bytes is smaller than a SceneHeader", bytes);
  CLRT_CUDA(cudaSetDevice(ctx->device));
  clrt_scene_header hdr; memcpy(&hdr, h_scene, sizeof hdr);
  if (hdr.totalDataSize != bytes) CLRT_FAIL("clrt_scene_upload: SceneHeader.totalDataSize (%llu) != buffer size (%zu)", (unsigned long long)hdr.totalDataSize, bytes);
  if (hdr.totalNumberOfTriangles >= (1ull << 30)) CLRT_FAIL("clrt_scene_upload: %llu triangles exceed 2^30-1", (unsigned long long)hdr.totalNumberOfTriangles);
  const uint8_t* h = (const uint8_t*)h_scene;
  // host walk of the headers, once (what getTriangleRefByIndex does per fetch)
  size_t off = sizeof(clrt_scene_header) + hdr.numberOfLights * 16 + hdr.numberOfSpheres * 16 + hdr.numberOfMaterials * sizeof(clrt_material);
  std::vector<MeshEntry> meshes;
  uint64_t tri = 0;
  for (uint64_t m = 0; m < hdr.numberOfModels; m++) {
    if (off + sizeof(clrt_model_header) > bytes) CLRT_FAIL("clrt_scene_upload: model %llu header runs past the buffer", (unsigned long long)m);
    clrt_model_header mh; memcpy(&mh, h + off, sizeof mh);
    size_t moff = off + sizeof(clrt_model_header);
    for (uint64_t s = 0; s < mh.numberOfSubmeshes; s++) {
      if (moff + sizeof(clrt_mesh_header) > bytes) CLRT_FAIL("clrt_scene_upload: model %llu sub-mesh %llu header runs past the buffer", (unsigned long long)m, (unsigned long long)s);
      clrt_mesh_header sh; memcpy(&sh, h + moff, sizeof sh);
      if (moff + sh.dataSize > bytes || sh.dataSize < sizeof(clrt_mesh_header) + sh.numberOfVertices * 16 + sh.numberOfIndices * 2)
        CLRT_FAIL("clrt_scene_upload: model %llu sub-mesh %llu is inconsistent (dataSize %llu)", (unsigned long long)m, (unsigned long long)s, (unsigned long long)sh.dataSize);
      MeshEntry e;
      e.byteOffset = moff; e.firstTriangle = (uint32_t)tri; e.numTriangles = (uint32_t)sh.numberOfTriangles;
      e.numVertices = (uint32_t)sh.numberOfVertices; e.materialIndex = (uint32_t)sh.materialIndex;
      e.modelIndex = (uint32_t)m; e.submeshIndex = (uint32_t)s;
      if (e.numTriangles) meshes.push_back(e);
      tri += sh.numberOfTriangles;
      moff += sh.dataSize;
    }
    off += mh.dataSize;
  }
  if (tri != hdr.totalNumberOfTriangles)
    CLRT_FAIL("clrt_scene_upload: sub-meshes hold %llu triangles, SceneHeader says %llu", (unsigned long long)tri, (unsigned long long)hdr.totalNumberOfTriangles);

  clrt_scene* s = new clrt_scene();
  s->ctx = ctx; s->bytes = bytes; s->header = hdr; s->meshes = meshes;
  s->n_tris = (uint32_t)tri; s->n_lights = (uint32_t)hdr.numberOfLights; s->n_materials = (uint32_t)hdr.numberOfMaterials;
  auto fail = [&](void) { clrt_scene_destroy(s); return 1; };
  if (cudaMalloc(&s->d_scene, bytes) != cudaSuccess) { clrt_set_error("clrt_scene_upload: cudaMalloc(%zu) failed", bytes); return fail(); }
  if (cudaMemcpyAsync(s->d_scene, h_scene, bytes, cudaMemcpyHostToDevice, ctx->stream) != cudaSuccess) { clrt_set_error("clrt_scene_upload: H2D copy failed"); return fail(); }
  const size_t nm = meshes.size() ? meshes.size() : 1;
  if (cudaMalloc(&s->d_meshes, nm * sizeof(MeshEntry)) != cudaSuccess) { clrt_set_error("clrt_scene_upload: cudaMalloc failed"); return fail(); }
  if (meshes.size()) cudaMemcpyAsync(s->d_meshes, meshes.data(), meshes.size() * sizeof(MeshEntry), cudaMemcpyHostToDevice, ctx->stream);
  const size_t nt = s->n_tris ? s->n_tris : 1;
  if (cudaMalloc(&s->d_tris, nt * sizeof(clrt_triangle)) != cudaSuccess || cudaMalloc(&s->d_triInMesh, nt * 4) != cudaSuccess) {
    clrt_set_error("clrt_scene_upload: cudaMalloc of the triangle array failed"); return fail(); }
  // aligned copies of lights and materials (they are 16-byte aligned in the packed buffer as well; copies keep
  // the shading kernels independent of the packed layout)
  const size_t nl = s->n_lights ? s->n_lights : 1, nmat = s->n_materials ? s->n_materials : 1;
  if (cudaMalloc(&s->d_lights, nl * sizeof(clrt_light)) != cudaSuccess || cudaMalloc(&s->d_materials, nmat * sizeof(clrt_material)) != cudaSuccess) {
    clrt_set_error("clrt_scene_upload: cudaMalloc failed"); return fail(); }
  const size_t lights_off = sizeof(clrt_scene_header);
  const size_t mats_off = lights_off + hdr.numberOfLights * 16 + hdr.numberOfSpheres * 16;
  if (s->n_lights) cudaMemcpyAsync(s->d_lights, h + lights_off, s->n_lights * sizeof(clrt_light), cudaMemcpyHostToDevice, ctx->stream);
  if (s->n_materials) cudaMemcpyAsync(s->d_materials, h + mats_off, s->n_materials * sizeof(clrt_material), cudaMemcpyHostToDevice, ctx->stream);
  if (s->n_tris) {
    flatten_triangles_kernel<<<(s->n_tris + 255) / 256, 256, 0, ctx->stream>>>(s->d_scene, s->d_meshes, (uint32_t)meshes.size(),
                                                                             s->n_tris, s->d_tris, s->d_triInMesh);
    if (cudaGetLastError() != cudaSuccess) { clrt_set_error("clrt_scene_upload: flatten kernel launch failed"); return fail(); }
  }
  cudaError_t e = cudaStreamSynchronize(ctx->stream);     // h_scene / meshes may go away after return
  if (e != cudaSuccess) { clrt_set_error("clrt_scene_upload: %s", cudaGetErrorString(e)); return fail(); }
  *out = s;
  return 0;
}

extern "C" int clrt_scene_destroy(clrt_scene* s) {
  if (!s) return 0;
  cudaSetDevice(s->ctx->device);
  cudaStreamSynchronize(s->ctx->stream);
  if (s->d_scene) cudaFree(s->d_scene);
  if (s->d_meshes) cudaFree(s->d_meshes);
  if (s->d_tris) cudaFree(s->d_tris);
  if (s->d_triInMesh) cudaFree(s->d_triInMesh);
  if (s->d_lights) cudaFree(s->d_lights);
  if (s->d_materials) cudaFree(s->d_materials);
  delete s;
  return 0;
}

extern "C" uint64_t clrt_scene_triangle_count(const clrt_scene* s) { return s ? s->n_tris : 0; }
extern "C" void* clrt_scene_dptr(const clrt_scene* s) { return s ? s->d_scene : nullptr; }
extern "C" void* clrt_scene_triangles_dptr(const clrt_scene* s) { return s ? s->d_tris : nullptr; }
extern "C" int clrt_scene_export_triangles(const clrt_scene* s, void* h_out) {
  if (!s || !h_out) CLRT_FAIL("clrt_scene_export_triangles: NULL argument");
  CLRT_CUDA(cudaSetDevice(s->ctx->device));
  CLRT_CUDA(cudaMemcpyAsync(h_out, s->d_tris, (size_t)s->n_tris * sizeof(clrt_triangle), cudaMemcpyDeviceToHost, s->ctx->stream));
  CLRT_CUDA(cudaStreamSynchronize(s->ctx->stream));
  return 0;
}
