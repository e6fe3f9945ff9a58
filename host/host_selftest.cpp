// host_selftest.cpp — exercises the C++ mirror (include/clrt/CLRayTracer.hpp) the way a caller of the reference would:
// building blocks with property checks of the reference's own test headers (R/Include/Testing/sortingTest.h,
// PrefixSumTest.h: sortedness + permutation, scan against a serial sum), then Scene -> BVHManager / TwoLevelGridManager ->
// generateContacts through the class interface, dumping the contacts for the Python parity tests.
//   host_selftest <scene.txt|packed.bin> <W> <H> <out_prefix> [bvh-only]
#include <clrt/CLRayTracer.hpp>

#include <cstdio>
#include <cstdlib>
#include <fstream>
#include <iostream>
#include <vector>

using namespace CLRayTracer;
using namespace CLRayTracer::Common;
using namespace CLRayTracer::OpenCLUtils;
using namespace CLRayTracer::AccelerationStructures;

#define REQUIRE(cond) if (!(cond)) { std::cout << "FAILED: " #cond " (line " << __LINE__ << ")\n" << err; return 1; }

static uint64_t splitmix64(uint64_t x) {
  x += 0x9E3779B97F4A7C15ull; x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull; x = (x ^ (x >> 27)) * 0x94D049BB133111EBull; return x ^ (x >> 31);
}
static bool dump(const std::string& path, const void* p, size_t bytes) {
  std::ofstream f(path.c_str(), std::ios::binary); f.write(static_cast<const char*>(p), (std::streamsize)bytes); return (bool)f;
}

int main(int argc, char** argv) {
  if (argc < 5) { std::cout << "usage: host_selftest <scene.txt|packed.bin> <W> <H> <out_prefix>\n"; return 2; }
  const std::string scenePath = argv[1], out = argv[4];
  const int W = atoi(argv[2]), H = atoi(argv[3]);
  const int structures = (argc > 5 && std::string(argv[5]) == "bvh-only") ? 1 : 2;
  Errata err;
  try {
    CLExecutionContext context(0);
    REQUIRE(context.getMaxComputeUnits() > 0 && context.getMaxWorkGroupSize() >= 256);

    // --- BitonicSort (Sorting.h:54-85): sorted by key, a permutation of the input, stable
    const size_t n = 1u << 16;
    std::vector<clrt_uint2> pairs(n), sorted(n);
    for (size_t i = 0; i < n; i++) { pairs[i].x = (uint32_t)(splitmix64(i) & 0x3FFF); pairs[i].y = (uint32_t)i; }
    {
      CLBuffer buf(context, n * sizeof(clrt_uint2), pairs.data(), CLBufferFlags::ReadWrite);
      BitonicSort sorter(context, true);
      REQUIRE(sorter.initialize(err) == Success);
      REQUIRE(sorter.sort(buf.getCLMem(), n, err) == Success);
      REQUIRE(buf.copyToHost(sorted.data(), err) == Success);
      std::vector<char> seen(n, 0);
      for (size_t i = 0; i < n; i++) {
        REQUIRE(sorted[i].y < n && !seen[sorted[i].y] && pairs[sorted[i].y].x == sorted[i].x);
        seen[sorted[i].y] = 1;
        if (i) REQUIRE(sorted[i - 1].x < sorted[i].x || (sorted[i - 1].x == sorted[i].x && sorted[i - 1].y < sorted[i].y));
      }
    }
    // --- PrefixSum (PrefixSum.h:48-83): inclusive
    {
      std::vector<uint32_t> in(n), res(n);
      for (size_t i = 0; i < n; i++) in[i] = (uint32_t)(splitmix64(i) & 7);
      CLBuffer a(context, n * 4, in.data(), CLBufferFlags::ReadOnly), b(context, n * 4, CLBufferFlags::ReadWrite);
      PrefixSum scan(context);
      REQUIRE(scan.initialize(err) == Success);
      REQUIRE(scan.computePrefixSum(a.getCLMem(), b.getCLMem(), n, err) == Success);
      REQUIRE(b.copyToHost(res.data(), err) == Success);
      uint32_t acc = 0;
      for (size_t i = 0; i < n; i++) { acc += in[i]; REQUIRE(res[i] == acc); }
    }
    // --- CLBuffer::resize is grow-only (CLBuffer.cpp:70-94)
    {
      CLBuffer b(context, 1024, CLBufferFlags::ReadWrite);
      b.resize(256); REQUIRE(b.getSize() == 256 && b.getActualSize() == 1024);
      b.resize(4096); REQUIRE(b.getSize() == 4096 && b.getActualSize() >= 4096);
    }
    // --- Scene + managers
    Scene scene;
    if (scenePath.size() > 4 && scenePath.substr(scenePath.size() - 4) == ".bin") {
      std::ifstream f(scenePath.c_str(), std::ios::binary);
      std::vector<char> bytes((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
      REQUIRE(scene.setHostSceneData(bytes.data(), bytes.size(), err) == Success);
    } else {
      REQUIRE(scene.load(scenePath.c_str(), err) == Success);
    }
    REQUIRE(dump(out + "_scene.bin", scene.getHostSceneData(), scene.getSceneDataSize()));
    {
      BVHManager unready(context, scene);
      Errata e2;
      if (unready.initialize(e2) != Error || e2.message().empty()) { std::cout << "FAILED: initialize before loadToGPU must report an error\n"; return 1; }
    }
    REQUIRE(scene.loadToGPU(context, err) == Success);
    const SceneHeader* sh = SCENE_HEADER(scene.getHostSceneData());
    Camera camera;
    camera.resX = W; camera.resY = H; camera.FOVDistance = FOVDistFromAngle(90, (float)W, (float)H); camera.supersampilngFactor = 1;
    CL_FLOAT3 pos;
    const float hx = (sh->modelsBoundingBox.bounds[1].x - sh->modelsBoundingBox.bounds[0].x) * 0.5f;
    const float hy = (sh->modelsBoundingBox.bounds[1].y - sh->modelsBoundingBox.bounds[0].y) * 0.5f;
    const float hz = (sh->modelsBoundingBox.bounds[1].z - sh->modelsBoundingBox.bounds[0].z) * 0.5f;
    pos.x = sh->modelsBoundingBox.bounds[0].x + hx; pos.y = sh->modelsBoundingBox.bounds[0].y + hy; pos.z = sh->modelsBoundingBox.bounds[0].z - hz; pos.w = 0;
    Quaternion q = zeroRotation();
    normalizeQuaternion(&q);
    setOrientationAndPos(&camera.viewTransform, q, pos);
    REQUIRE(dump(out + "_camera.bin", &camera, sizeof(camera)));
    std::vector<Contact> contacts((size_t)W * H);
    for (int which = 0; which < structures; which++) {
      std::shared_ptr<AccelerationStructureManager> m;
      if (which == 0) m.reset(new BVHManager(context, scene)); else m.reset(new TwoLevelGridManager(context, scene));
      REQUIRE(m->initialize(err) == Success);
      REQUIRE(m->initializeFrame(err) == Success);
      REQUIRE(m->construct(err) == Success);
      if (which == 0) REQUIRE(m->setTraceMode(CLRT_TRACE_REFERENCE_ORDER, err) == Success);
      REQUIRE(m->generateContacts(camera, err) == Success);
      REQUIRE(m->getPrimaryContacts()->getActualSize() / sizeof(Contact) == contacts.size());
      REQUIRE(m->getPrimaryContacts()->copyToHost(contacts.data(), err) == Success);
      REQUIRE(dump(out + (which == 0 ? "_bvh_contacts.bin" : "_tlg_contacts.bin"), contacts.data(), contacts.size() * sizeof(Contact)));
      // generic-ray entry point: the same rays through generateContacts(rays, contacts, n) give the same contacts
      CLBuffer rays(context, contacts.size() * sizeof(Ray), CLBufferFlags::ReadWrite), out2(context, contacts.size() * sizeof(Contact), CLBufferFlags::ReadWrite);
      REQUIRE(clrt_generate_primary_rays(context.handle(), &camera, rays.getCLMem()) == 0);
      REQUIRE(m->generateContacts(rays, out2, (unsigned)contacts.size(), err) == Success);
      std::vector<Contact> c2(contacts.size());
      REQUIRE(out2.copyToHost(c2.data(), err) == Success);
      size_t same = 0;
      for (size_t i = 0; i < c2.size(); i++)
        same += c2[i].normalAndintersectionDistance.w == contacts[i].normalAndintersectionDistance.w && c2[i].pad[0] == contacts[i].pad[0];
      REQUIRE(same == c2.size());
      size_t hits = 0;
      for (size_t i = 0; i < contacts.size(); i++) hits += contacts[i].normalAndintersectionDistance.w > 0;
      std::cout << (which == 0 ? "BVH" : "GRID") << ": " << hits << " of " << contacts.size() << " rays hit\n";
    }
  } catch (const CLInterfaceException& e) {
    std::cout << "exception: " << static_cast<const Errata&>(e);
    return 1;
  }
  std::cout << "host_selftest ok\n";
  return 0;
}
