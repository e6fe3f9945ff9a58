// oracle/ref_bridge.cpp — TEST INFRASTRUCTURE, never part of the product path.
//
// Compiles the reference's OWN dual-target headers (R/Include/CLData/**, where they lie under
// /root/reference; nothing is copied into this repository) on the host, and exposes the
// reference arithmetic through a small C ABI -> oracle/_ref/libclrt_ref.so (git-ignored, travels
// to the GPU box).  It is used (a) to pin oracle/clrt_oracle.cpp (the CPU restatement) and
// (b) as the "reference" CPU baseline in bench.py.
//
// R = /root/reference/OpenCLRayTracer.  What is the reference's code and what is restated here:
//   reference code, called as-is : calculateMorton (R/Include/CLData/AccelerationStructs/BVH.h:83-130),
//        constructNode (:291-321), mergeBoundingBox (:328-334), bvh_generate_contact (:344-408),
//        generateRay (R/Include/CLData/CLStructs.h:95-106), cpuShadeBlinnPhong (R/Include/CLData/Shading.h:74-166),
//        the Two-Level-Grid functions (R/Include/CLData/AccelerationStructs/TwoLevelGrid.h), tinyobj::LoadObj
//        (R/Include/3rdParty/tiny_obj_loader.h), and every struct layout.
//   restated here (host orchestration that needs Windows/boost/OpenCL and cannot be compiled):
//        Scene::load packing loop (R/src/CLRayTracerEngine/Scene.cpp:231-444), the kernel launch loops
//        (R/src/CLRayTracerEngine/BVHKernels.cl:46-117, BVHManager.cpp:122-228), the demo camera rule
//        (R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp:202-217) and shading loop (:301-316).
//
// Host-vs-device gaps closed by hand: `sign` macro (Q22) is replaced by OpenCL sign();
// vector types are 4-byte aligned PODs (Q14); struct Ray is converted from the 48-byte device layout.
#include <stdint.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <float.h>
#include <limits.h>
#include <ctime>
#include <vector>
#include <string>
#include <map>
#include <fstream>
#include <iostream>
#include <sstream>
#include <algorithm>
#ifdef _OPENMP
#include <omp.h>
#endif

#define TINYOBJLOADER_IMPLEMENTATION
#include <3rdParty/tiny_obj_loader.h>

// ---- enter "host build of the dual-target headers" mode --------------------------------------
#define _WIN32 1
#define __declspec(x) /* no over-alignment on the host: see Q14 */
#include <CLData/CLPortability.h>
// Q22: the host `sign` macro (CLPortability.h:282) has a precedence bug; the device runs OpenCL sign().
#undef sign
static inline float clrt_cl_sign(float v) { return v > 0.0f ? 1.0f : (v < 0.0f ? -1.0f : (v == 0.0f ? v : 0.0f)); }
#define sign(n) clrt_cl_sign(n)
#include <CLData/AccelerationStructs/BVH.h>
// Q20: `pow` in calcLeafCellResolution (TwoLevelGrid.h:199-205) is implementation defined on the OpenCL device and its result
// is truncated to a cell count; both sides of the parity check evaluate it in double and round to float once.
static inline float clrt_ref_pow(float x, float y) { return (float)::pow((double)x, (double)y); }
#define pow(x, y) clrt_ref_pow((x), (y))
#include <CLData/AccelerationStructs/TwoLevelGrid.h>
#undef pow
#include <CLData/Shading.h>
#undef _WIN32
// ----------------------------------------------------------------------------------------------

extern "C" {

// Struct sizes as compiled here; tests compare them with include/clrt.h's layouts.
void ref_struct_sizes(uint32_t* out) {
  out[0] = sizeof(struct BVHNode);   out[1] = sizeof(struct AABB);       out[2] = sizeof(struct Contact);
  out[3] = sizeof(struct Camera);    out[4] = sizeof(struct Matrix4);    out[5] = sizeof(struct SceneHeader);
  out[6] = sizeof(struct ModelHeader); out[7] = sizeof(struct MeshHeader); out[8] = sizeof(struct Material);
  out[9] = sizeof(struct Light);     out[10] = sizeof(struct GridData);  out[11] = sizeof(struct TopLevelCell);
  out[12] = sizeof(cl_float3);       out[13] = sizeof(cl_uint2);
}

// ---- Scene packing: restates Scene::load (Scene.cpp:231-444) with the reference's tinyobj -------
struct RefModel { std::vector<tinyobj::shape_t> shapes; std::vector<tinyobj::material_t> materials; unsigned long size; };

static Material ref_fill_material(tinyobj::material_t& m) {        // Scene.cpp:97-126
  Material r; memset(&r, 0, sizeof r);
  r.ambient.x = m.ambient[0]; r.ambient.y = m.ambient[1]; r.ambient.z = m.ambient[2];
  r.diffuse.x = m.diffuse[0]; r.diffuse.y = m.diffuse[1]; r.diffuse.z = m.diffuse[2];
  r.specular.x = m.specular[0]; r.specular.y = m.specular[1]; r.specular.z = m.specular[2];
  r.emission.x = m.emission[0]; r.emission.y = m.emission[1]; r.emission.z = m.emission[2];
  r.transmittance.x = m.transmittance[0]; r.transmittance.y = m.transmittance[1]; r.transmittance.z = m.transmittance[2];
  r.dissolve = m.dissolve; r.illum = m.illum; r.ior = m.ior; r.shininess = m.shininess;
  return r;
}

// lights: nLights x (x,y,z,energy).  objPaths: nObjs OBJ files.  Returns malloc'd scene buffer.
int ref_scene_pack(const float* lights, uint32_t nLights, const char** objPaths, uint32_t nObjs,
                   uint8_t** outBuf, uint64_t* outSize) {
  std::vector<RefModel> models; std::vector<Material> uniq;
  uint64_t total = sizeof(struct SceneHeader) + (uint64_t)nLights * sizeof(struct Light);
  uint64_t modelBytes = 0;
  for (uint32_t i = 0; i < nObjs; i++) {
    RefModel m; std::string err, file(objPaths[i]);
    size_t s = file.find_last_of('/');
    std::string base = (s == std::string::npos) ? std::string("") : file.substr(0, s + 1);
    if (!tinyobj::LoadObj(m.shapes, m.materials, err, file.c_str(), base.c_str())) return 1;
    unsigned long b = 0;                                             // Scene.cpp:80-95
    for (size_t k = 0; k < m.shapes.size(); k++)
      b += m.shapes[k].mesh.indices.size() * sizeof(INDEX_TYPE) + m.shapes[k].mesh.positions.size() / 3 * sizeof(VERTEX_TYPE) + MESH_HEADER_SIZE;
    m.size = b + MODEL_HEADER_SIZE; modelBytes += m.size;
    std::map<cl_uint, cl_uint> xf;                                   // Scene.cpp:129-167
    for (cl_uint j = 0; j < m.materials.size(); j++) {
      Material mat = ref_fill_material(m.materials[j]);
      bool found = false;
      for (cl_uint k = 0; k < uniq.size(); k++) if (materialEquals(&mat, &uniq[k])) { xf[j] = k; found = true; break; }
      if (!found) { xf[j] = (cl_uint)uniq.size(); uniq.push_back(mat); }
    }
    for (size_t k = 0; k < m.shapes.size(); k++)
      for (size_t j = 0; j < m.shapes[k].mesh.material_ids.size(); j++)
        m.shapes[k].mesh.material_ids[j] = xf[m.shapes[k].mesh.material_ids[j]];
    models.push_back(m);
  }
  total += modelBytes + uniq.size() * sizeof(struct Material);
  char* buf = (char*)calloc(total, 1);
  SceneHeader* sh = SCENE_HEADER(buf);
  sh->totalDataSize = total; sh->numberOfPrimitives = nLights + nObjs; sh->numberOfLights = nLights;
  sh->numberOfSpheres = 0; sh->numberOfModels = models.size(); sh->modelBufferSize = modelBytes;
  sh->numberOfMaterials = uniq.size();
  for (uint32_t i = 0; i < nLights; i++) {
    Light* lt = getLightAtIndex(buf, i);
    lt->posAndEnergy.x = lights[4*i]; lt->posAndEnergy.y = lights[4*i+1]; lt->posAndEnergy.z = lights[4*i+2]; lt->posAndEnergy.w = lights[4*i+3];
  }
  for (size_t i = 0; i < uniq.size(); i++) *getMaterialAtIndex(buf, (int)i) = uniq[i];
  fillVector3(sh->modelsBoundingBox.bounds[0], FLT_MAX, FLT_MAX, FLT_MAX);
  fillVector3(sh->modelsBoundingBox.bounds[1], FLT_MIN, FLT_MIN, FLT_MIN);
  CL_ULONG sceneTris = 0;
  for (size_t i = 0; i < models.size(); i++) {                       // Scene.cpp:390-440
    char* md = getModelAtIndex((CL_UINT)i, buf);
    ModelHeader* hdr = MODEL_HEADER(md);
    hdr->dataSize = models[i].size; hdr->numberOfSubmeshes = models[i].shapes.size();
    CL_FLOAT3 mn, mx; memset(&mn, 0, sizeof mn); memset(&mx, 0, sizeof mx);
    fillVector3(mn, FLT_MAX, FLT_MAX, FLT_MAX); fillVector3(mx, FLT_MIN, FLT_MIN, FLT_MIN);
    CL_ULONG modelTris = 0;
    for (size_t m = 0; m < models[i].shapes.size(); m++) {
      tinyobj::shape_t* shape = &models[i].shapes[m];
      // NB: the mesh header must be written before getMeshAtIndex can walk past it.
      char* mb = getMeshAtIndex((CL_UINT)m, md);
      const unsigned nv = shape->mesh.positions.size() / 3, ni = shape->mesh.indices.size();
      MeshHeader* mh = MESH_HEADER(mb);
      mh->dataSize = ni * sizeof(INDEX_TYPE) + nv * sizeof(VERTEX_TYPE) + MESH_HEADER_SIZE;
      mh->numberOfVertices = nv; mh->numberOfIndices = ni;
      mh->materialIndex = shape->mesh.material_ids.empty() ? 0 : shape->mesh.material_ids[0];
      mh->numberOfTriangles = ni / 3; modelTris += ni / 3;
      int p = 0;
      for (unsigned vi = 0; vi < nv; vi++) {
        VERTEX_TYPE v; v.w = 0.0f;                                   // the reference leaves .w indeterminate
        v.x = shape->mesh.positions[p++]; v.y = shape->mesh.positions[p++]; v.z = shape->mesh.positions[p++];
        mn.x = min(mn.x, v.x); mn.y = min(mn.y, v.y); mn.z = min(mn.z, v.z);
        mx.x = max(mx.x, v.x); mx.y = max(mx.y, v.y); mx.z = max(mx.z, v.z);
        setVertexAt(v, vi, mb);
      }
      for (unsigned ii = 0; ii < ni; ii++) setIndexAt((INDEX_TYPE)shape->mesh.indices[ii], ii, mb);
      hdr->boundingBox.bounds[0] = mn; hdr->boundingBox.bounds[1] = mx;
      sh->modelsBoundingBox = merge(hdr->boundingBox, sh->modelsBoundingBox);
    }
    hdr->numberOfTriangles = modelTris; sceneTris += modelTris;
  }
  sh->totalNumberOfTriangles = sceneTris;
  *outBuf = (uint8_t*)buf; *outSize = total;
  return 0;
}
void ref_free(void* p) { free(p); }

// ---- LBVH build ---------------------------------------------------------------------------------
// BVHKernels.cl:46-50 over all triangles.  nodes: (2N-1) x 48 B, pairs: N x 8 B.
void ref_bvh_morton(const uint8_t* scene, uint8_t* nodes, uint32_t* pairs) {
  const long n = (long)SCENE_HEADER((const char*)scene)->totalNumberOfTriangles;
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < n; i++) calculateMorton((struct BVHNode*)nodes, (CL_UINT2*)pairs, (CL_UINT)i, (const char*)scene);
}
// BVHKernels.cl:55-60.  literal=1 keeps the `idx < leafCount-2` guard (Q1); 0 builds all N-1 nodes.
void ref_bvh_radix_tree(uint8_t* nodes, uint32_t* sortedPairs, uint32_t n, int literal) {
  const long cnt = literal ? (long)n - 2 : (long)n - 1;
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < cnt; i++) constructNode((struct BVHNode*)nodes, (CL_UINT2*)sortedPairs, n, (int)i);
}
// BVHKernels.cl:65-79, work-items run one after another (the final boxes are order independent: exact min/max).
void ref_bvh_refit(uint8_t* nodesRaw, uint32_t n) {
  struct BVHNode* nodes = (struct BVHNode*)nodesRaw;
  std::vector<uint32_t> counters(n, 0u);
  for (uint32_t gid = 0; gid < n; gid++) {
    CL_UINT current = parent(nodes[gid]);
    while (current != UINT_MAX) {
      counters[current - n]++;
      if (counters[current - n] > 1) { mergeBoundingBox(nodes, current); current = parent(nodes[current]); }
      else break;
    }
  }
}

// ---- Tracing ------------------------------------------------------------------------------------
// BVHKernels.cl:84-98.  contacts: resX*resY x 32 B.
void ref_bvh_trace_primary(const uint8_t* camera, const uint8_t* nodes, uint32_t n, const uint8_t* scene, uint8_t* contacts) {
  struct Camera cam; memcpy(&cam, camera, sizeof cam);
  const long rays = (long)cam.resX * cam.resY;
  #pragma omp parallel for schedule(dynamic, 256)
  for (long i = 0; i < rays; i++) {
    struct Ray r = generateRay(&cam, (CL_UINT)i);
    struct Contact c = bvh_generate_contact(r, (struct BVHNode*)nodes, n, (const char*)scene);
    c.pixelIndex = (CL_UINT)i;
    memcpy(contacts + 32 * i, &c, 32);
  }
}
// BVHKernels.cl:103-117.  rays use the 48-byte device layout {uint idx; pad[3]; float4 origin; float4 direction}.
void ref_bvh_trace_rays(const uint8_t* rays48, uint32_t rayCount, const uint8_t* nodes, uint32_t n, const uint8_t* scene, uint8_t* contacts) {
  #pragma omp parallel for schedule(dynamic, 256)
  for (long i = 0; i < (long)rayCount; i++) {
    struct Ray r; memset(&r, 0, sizeof r);
    memcpy(&r.idx, rays48 + 48 * i, 4); memcpy(&r.origin, rays48 + 48 * i + 16, 16); memcpy(&r.direction, rays48 + 48 * i + 32, 16);
    struct Contact c = bvh_generate_contact(r, (struct BVHNode*)nodes, n, (const char*)scene);
    c.pixelIndex = (CL_UINT)i;
    memcpy(contacts + 32 * i, &c, 32);
  }
}
// generateRay only (CLStructs.h:95-106): out = rays x 48 B device layout.
void ref_generate_rays(const uint8_t* camera, uint8_t* rays48) {
  struct Camera cam; memcpy(&cam, camera, sizeof cam);
  const long rays = (long)cam.resX * cam.resY;
  for (long i = 0; i < rays; i++) {
    struct Ray r = generateRay(&cam, (CL_UINT)i);
    memset(rays48 + 48 * i, 0, 48);
    memcpy(rays48 + 48 * i, &r.idx, 4); memcpy(rays48 + 48 * i + 16, &r.origin, 12); memcpy(rays48 + 48 * i + 32, &r.direction, 12);
  }
}

// ---- Demo camera rule + host shading loop -------------------------------------------------------
// OpenCLRayTracerDemo.cpp:202-217 then RTFrame :275-276 with the zero-rotation quaternion.
void ref_demo_camera(const uint8_t* scene, uint32_t w, uint32_t h, uint8_t* cameraOut) {
  const SceneHeader* sh = SCENE_HEADER((const char*)scene);
  struct Camera cam; memset(&cam, 0, sizeof cam);
  cam.resX = w; cam.resY = h; cam.FOVDistance = FOVDistFromAngle(90, (float)w, (float)h); cam.supersampilngFactor = 1;
  CL_FLOAT3 half, pos; memset(&pos, 0, sizeof pos);
  half.x = (sh->modelsBoundingBox.bounds[1].x - sh->modelsBoundingBox.bounds[0].x) * 0.5f;
  half.y = (sh->modelsBoundingBox.bounds[1].y - sh->modelsBoundingBox.bounds[0].y) * 0.5f;
  half.z = (sh->modelsBoundingBox.bounds[1].z - sh->modelsBoundingBox.bounds[0].z) * 0.5f;
  pos.x = sh->modelsBoundingBox.bounds[0].x + half.x;
  pos.y = sh->modelsBoundingBox.bounds[0].y + half.y;
  pos.z = sh->modelsBoundingBox.bounds[0].z - half.z;
  cam.viewTransform = identityTransform();
  struct Quaternion q = zeroRotation();
  normalizeQuaternion(&q);
  setOrientationAndPos(&cam.viewTransform, q, pos);
  memcpy(cameraOut, &cam, sizeof cam);
}
// General camera: quaternion (r,i,j,k) + position, as RTFrame does each frame.
void ref_camera_from_pose(uint32_t w, uint32_t h, float fovDeg, const float* quat, const float* pos3, uint8_t* cameraOut) {
  struct Camera cam; memset(&cam, 0, sizeof cam);
  cam.resX = w; cam.resY = h; cam.FOVDistance = FOVDistFromAngle(fovDeg, (float)w, (float)h); cam.supersampilngFactor = 1;
  struct Quaternion q; q.r = quat[0]; q.i = quat[1]; q.j = quat[2]; q.k = quat[3];
  CL_FLOAT3 pos; memset(&pos, 0, sizeof pos); pos.x = pos3[0]; pos.y = pos3[1]; pos.z = pos3[2];
  cam.viewTransform = identityTransform();
  normalizeQuaternion(&q);
  setOrientationAndPos(&cam.viewTransform, q, pos);
  memcpy(cameraOut, &cam, sizeof cam);
}
// OpenCLRayTracerDemo.cpp:298-316: colours float4 per pixel, zero where contactDist <= 0.
void ref_shade(const uint8_t* camera, const uint8_t* contacts, uint32_t count, const uint8_t* scene, float* colors4) {
  struct Camera cam; memcpy(&cam, camera, sizeof cam);
  memset(colors4, 0, (size_t)count * 16);
  #pragma omp parallel for schedule(static)
  for (long i = 0; i < (long)count; i++) {
    struct Contact c; memcpy(&c, contacts + 32 * i, 32);
    if (c.contactDist > 0) {
      CL_FLOAT4 color; color.x = color.y = color.z = 0.0f; color.w = 1.0f;
      cpuShadeBlinnPhong(&cam, &c, (char*)scene, color);
      memcpy(colors4 + 4 * (size_t)c.pixelIndex, &color, 16);
    }
  }
}

// ---- Two-Level Grid ---------------------------------------------------------------------------------------------
// Reference code called as-is: prepareGridData (TwoLevelGrid.h:102-124), writePairs (:173-190), fillTopLevelCell
// (:215-225), countLeafPairs (:295-315), writeLeafPairs (:378-403), tlg_generate_contact (:557-678).
// Restated here (host code that needs OpenCL/boost): TwoLevelGridManager::getResolution/calculateGridData
// (TwoLevelGridManager.cpp:566-600), the construct() sequence (:207-469), the scan (PrefixSum.cpp, inclusive) and the two
// range-extraction kernels (TwoLevelGridKernels.cl:71-116,179-221) with the FIXED semantics of SURVEY.md Q17/Q18/Q21:
// every triangle takes part; range[c] = [first, one-past-last) over the REAL pairs; buffers zero-initialised; the sort is
// the canonical stable one.  literal != 0 keeps Q17 (the last triangle is skipped by count and scatter).
struct RefTlg {
  GridData grid;
  std::vector<TopLevelCell> cells;
  std::vector<CL_UINT2> topRanges, topPairs, leafRanges, leafPairs;   // pairs: real ones only, sorted
  uint32_t leafCells;
};

static void ref_extract_ranges(const std::vector<CL_UINT2>& sorted, std::vector<CL_UINT2>& ranges) {
  const size_t n = sorted.size();
  for (size_t i = 0; i + 1 < n; i++) {
    const CL_UINT thisOne = sorted[i].x, right = sorted[i + 1].x;
    if (right > thisOne) { ranges[thisOne].y = (CL_UINT)(i + 1); ranges[right].x = (CL_UINT)(i + 1); }
  }
  if (n) ranges[sorted[n - 1].x].y = (CL_UINT)n;                       // FIXED: one past the last (the reference writes n-1)
}
static void ref_stable_sort(std::vector<CL_UINT2>& v) {
  std::stable_sort(v.begin(), v.end(), [](const CL_UINT2& a, const CL_UINT2& b) { return a.x < b.x; });
}

void* ref_tlg_build(const uint8_t* sceneRaw, float topLevelDensity, float leafDensity, int literal) {
  const char* scene = (const char*)sceneRaw;
  const SceneHeader* sh = SCENE_HEADER(scene);
  RefTlg* t = new RefTlg();
  // getResolution + calculateGridData (TwoLevelGridManager.cpp:566-600)
  {
    struct AABB bounds = sh->modelsBoundingBox;
    float deltaX = bounds.bounds[1].x - bounds.bounds[0].x;
    float deltaY = bounds.bounds[1].y - bounds.bounds[0].y;
    float deltaZ = bounds.bounds[1].z - bounds.bounds[0].z;
    float volume = boxVolume(bounds);
    float prims = sh->totalNumberOfTriangles;
    float a = clrt_ref_pow(topLevelDensity * prims / volume, oneThird);
    memset(&t->grid, 0, sizeof t->grid);
    t->grid.box = bounds;
    t->grid.resX = deltaX * a; t->grid.resY = deltaY * a; t->grid.resZ = deltaZ * a;
    t->grid.stepX = (bounds.bounds[1].x - bounds.bounds[0].x) / t->grid.resX;
    t->grid.stepY = (bounds.bounds[1].y - bounds.bounds[0].y) / t->grid.resY;
    t->grid.stepZ = (bounds.bounds[1].z - bounds.bounds[0].z) / t->grid.resZ;
    t->grid.leafDensity = leafDensity;
  }
  const uint32_t n = (uint32_t)sh->totalNumberOfTriangles;
  const uint32_t cellsCount = t->grid.resX * t->grid.resY * t->grid.resZ;
  if (cellsCount == 0) { t->leafCells = 0; return t; }
  const uint32_t tris = literal ? n - 1 : n;                                 // Q17
  std::vector<CL_UINT> counters(n, 0), scan(n, 0);
  for (uint32_t i = 0; i < tris; i++) prepareGridData(scene, i, &t->grid, counters.data());       // TwoLevelGridKernels.cl:43-51
  { CL_UINT run = 0; for (uint32_t i = 0; i < n; i++) { run += counters[i]; scan[i] = run; } }   // inclusive (Q23)
  const uint32_t pairsCount = n ? scan[n - 1] : 0;
  t->topPairs.assign(pairsCount, CL_UINT2());
  for (auto& p : t->topPairs) p.x = p.y = UINT_MAX;
  for (uint32_t i = 0; i < tris; i++) writePairs(scene, i, &t->grid, scan.data(), counters.data(), t->topPairs.data());   // :56-66
  ref_stable_sort(t->topPairs);
  t->topRanges.assign(cellsCount, CL_UINT2()); for (auto& r : t->topRanges) r.x = r.y = 0;
  ref_extract_ranges(t->topPairs, t->topRanges);
  t->cells.assign(cellsCount, TopLevelCell()); memset(t->cells.data(), 0, cellsCount * sizeof(TopLevelCell));
  std::vector<CL_UINT> leafCounts(cellsCount, 0), leafScan(cellsCount, 0);
  for (uint32_t c = 0; c < cellsCount; c++) fillTopLevelCell(t->topRanges.data(), leafCounts.data(), t->cells.data(), &t->grid, c);   // :121-127
  { CL_UINT run = 0; for (uint32_t c = 0; c < cellsCount; c++) { run += leafCounts[c]; leafScan[c] = run; } }
  t->leafCells = leafScan[cellsCount - 1];
  for (uint32_t c = 1; c < cellsCount; c++) t->cells[c].firstLeafIdx = leafScan[c - 1];        // :132-139
  std::vector<CL_UINT> lcount(pairsCount, 0), lscan(pairsCount, 0);
  #pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < (long)pairsCount; p++)
    lcount[p] = countLeafPairs(scene, t->topPairs.data(), (CL_UINT)p, &t->grid, t->cells.data());                         // :144-157
  { CL_UINT run = 0; for (uint32_t p = 0; p < pairsCount; p++) { run += lcount[p]; lscan[p] = run; } }
  const uint32_t leafPairsCap = pairsCount ? lscan[pairsCount - 1] : 0;
  std::vector<CL_UINT2> lp(leafPairsCap);
  for (auto& p : lp) p.x = p.y = UINT_MAX;
  #pragma omp parallel for schedule(dynamic, 64)
  for (long p = 0; p < (long)pairsCount; p++)
    writeLeafPairs(scene, t->topPairs.data(), t->cells.data(), (CL_UINT)p, &t->grid, lscan.data(), lcount.data(), lp.data());   // :162-174
  ref_stable_sort(lp);
  size_t real = 0; while (real < lp.size() && lp[real].x != UINT_MAX) real++;    // Q19: gaps (UINT_MAX) sort last
  lp.resize(real);
  t->leafPairs.swap(lp);
  t->leafRanges.assign(t->leafCells, CL_UINT2()); for (auto& r : t->leafRanges) r.x = r.y = 0;   // Q21: cleared
  ref_extract_ranges(t->leafPairs, t->leafRanges);
  return t;
}
void ref_tlg_free(void* h) { delete (RefTlg*)h; }
void ref_tlg_sizes(void* h, uint32_t* out4) {
  RefTlg* t = (RefTlg*)h;
  out4[0] = (uint32_t)t->cells.size(); out4[1] = (uint32_t)t->topPairs.size(); out4[2] = t->leafCells; out4[3] = (uint32_t)t->leafPairs.size();
}
void ref_tlg_export(void* h, uint8_t* grid, uint8_t* cells, uint8_t* topRanges, uint8_t* topPairs, uint8_t* leafRanges, uint8_t* leafPairs) {
  RefTlg* t = (RefTlg*)h;
  memcpy(grid, &t->grid, sizeof(GridData));
  if (t->cells.size()) { memcpy(cells, t->cells.data(), t->cells.size() * 16); memcpy(topRanges, t->topRanges.data(), t->topRanges.size() * 8); }
  if (t->topPairs.size()) memcpy(topPairs, t->topPairs.data(), t->topPairs.size() * 8);
  if (t->leafRanges.size()) memcpy(leafRanges, t->leafRanges.data(), t->leafRanges.size() * 8);
  if (t->leafPairs.size()) memcpy(leafPairs, t->leafPairs.data(), t->leafPairs.size() * 8);
}
// TwoLevelGridKernels.cl:228-246 / :251-267
void ref_tlg_trace_primary(void* h, const uint8_t* camera, const uint8_t* scene, uint8_t* contacts) {
  RefTlg* t = (RefTlg*)h;
  struct Camera cam; memcpy(&cam, camera, sizeof cam);
  const long rays = (long)cam.resX * cam.resY;
  #pragma omp parallel for schedule(dynamic, 256)
  for (long i = 0; i < rays; i++) {
    const struct Ray r = generateRay(&cam, (CL_UINT)i);
    struct Contact c = tlg_generate_contact(r, (const char*)scene, &t->grid, t->cells.data(), t->leafRanges.data(), t->leafPairs.data());
    c.pixelIndex = (CL_UINT)i;
    memcpy(contacts + 32 * i, &c, 32);
  }
}
void ref_tlg_trace_rays(void* h, const uint8_t* rays48, uint32_t rayCount, const uint8_t* scene, uint8_t* contacts) {
  RefTlg* t = (RefTlg*)h;
  #pragma omp parallel for schedule(dynamic, 256)
  for (long i = 0; i < (long)rayCount; i++) {
    struct Ray r; memset(&r, 0, sizeof r);
    memcpy(&r.idx, rays48 + 48 * i, 4); memcpy(&r.origin, rays48 + 48 * i + 16, 16); memcpy(&r.direction, rays48 + 48 * i + 32, 16);
    struct Contact c = tlg_generate_contact(r, (const char*)scene, &t->grid, t->cells.data(), t->leafRanges.data(), t->leafPairs.data());
    c.pixelIndex = (CL_UINT)i;
    memcpy(contacts + 32 * i, &c, 32);
  }
}

// torchrun exports OMP_NUM_THREADS=1 to its workers; the CPU baseline is specified on all host cores (bench.py sets the count).
void ref_set_num_threads(int n) {
#ifdef _OPENMP
  if (n > 0) omp_set_num_threads(n);
#else
  (void)n;
#endif
}

int ref_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}

} // extern "C"
