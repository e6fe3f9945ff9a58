// OpenCLRayTracerDemo.cpp — the reference's demo entry point, headless.
//
// Follows R/src/Demos/OpenCLRayTracerDemo/OpenCLRayTracerDemo.cpp (R = /root/reference/OpenCLRayTracer): the same switches
// (:66-73,513-552: -winH -winW -accStruct {BVH,GRID} -scene; -headersPath is accepted and ignored, nothing is compiled at
// run time), the same start-up sequence (:131-239: load scene, loadToGPU, demo camera, manager initialize / initializeFrame /
// construct) and the same frame (:271-326: setOrientationAndPos -> generateContacts(camera) -> shading).  What cannot exist
// here: the GLUT window and CL/GL interop (graphics stack: unavailable on this box).  The window becomes a device
// framebuffer written to a binary PPM, and the keyboard / mouse callbacks (:378-451) become a script:
//   -script "UUDLRadws"   one frame per letter: U/D/L/R = arrow keys (move 0.3 along forward / side), a/d = mouse drag
//                         left/right, w/s = drag up/down (5 degrees per event)
//   -scene x.bin          a packed scene buffer instead of a scene file (Scene::setHostSceneData)
//   -out prefix           writes prefix_<frame>.ppm (default: no files);  -shadows adds the shadow pass;  -device N
// Shading runs on the device (clrt_shade; the reference shades on the host with the same Blinn-Phong, Shading.h:74-166).
#include <clrt/CLRayTracer.hpp>

#include <cstdio>
#include <cstdlib>
#include <iostream>
#include <vector>

using namespace CLRayTracer;
using namespace CLRayTracer::Common;
using namespace CLRayTracer::OpenCLUtils;
using namespace CLRayTracer::AccelerationStructures;

enum AccelerationStruct { INVALID = -1, BVH = 0, GRID = 1, KD = 2 };       // KD: ours (the reference's TODO 2.1)

static void printUsage() {
  std::cout << "OpenCLRayTracer demo (headless, CUDA/B200 back end)\nCommand line parameters:\n"
            << "-winH <Window Height>\n-winW <Window Width>\n-accStruct <Acceleration Structure> - valid values: BVH, GRID (and KD, not in the reference)\n"
            << "-scene <Path to scene file>\n[-headersPath <ignored>] [-script <UDLRadws...>] [-out <ppm prefix>] [-shadows] [-device <n>]\n";
}

#define CHECKED_CALL(result) if ((result) != Success) { std::cout << err; return 1; }

static bool write_ppm(const std::string& path, const std::vector<unsigned char>& rgba, int w, int h) {
  FILE* f = fopen(path.c_str(), "wb");
  if (!f) return false;
  fprintf(f, "P6\n%d %d\n255\n", w, h);
  std::vector<unsigned char> row((size_t)w * 3);
  for (int y = 0; y < h; y++) {
    for (int x = 0; x < w; x++) for (int c = 0; c < 3; c++) row[(size_t)x * 3 + c] = rgba[((size_t)y * w + x) * 4 + c];
    fwrite(row.data(), 1, row.size(), f);
  }
  fclose(f);
  return true;
}

int main(int argc, char* argv[]) {
  int window_width = 0, window_height = 0, device = 0;
  AccelerationStruct accelerationStructInUse = INVALID;
  std::string scenePath, script, outPrefix;
  bool shadows = false;
  for (int i = 0; i < argc; i++) {
    const std::string current = argv[i];
    const char* next = i + 1 < argc ? argv[i + 1] : NULL;
    if (current == "-winH") window_height = next ? atoi(next) : 0;
    else if (current == "-winW") window_width = next ? atoi(next) : 0;
    else if (current == "-accStruct" && next) accelerationStructInUse = std::string(next) == "BVH" ? BVH : std::string(next) == "GRID" ? GRID : std::string(next) == "KD" ? KD : INVALID;
    else if (current == "-scene") scenePath = next ? next : "";
    else if (current == "-script") script = next ? next : "";
    else if (current == "-out") outPrefix = next ? next : "";
    else if (current == "-device") device = next ? atoi(next) : 0;
    else if (current == "-shadows") shadows = true;
  }
  if (!(window_height > 0 && window_width > 0 && accelerationStructInUse != INVALID && !scenePath.empty())) { printUsage(); return 0; }
  const size_t pixelCount = (size_t)window_width * window_height;
  std::cout << "Configuration parameters:\n    Window width: " << window_width << "\n    Window heigth: " << window_height
            << "\n    Acceleration struct: " << (accelerationStructInUse == BVH ? "BVH" : accelerationStructInUse == GRID ? "GRID" : "KD") << "\n    Scene file: " << scenePath << std::endl;

  Errata err;
  try {
    CLExecutionContext context(device);
    std::cout << "Loading scene: " << scenePath << std::endl;
    Scene scene;
    if (scenePath.size() > 4 && scenePath.substr(scenePath.size() - 4) == ".bin") {
      // an already packed v0.6 scene buffer (what Scene::load produces; synthetic scenes and the test fixtures come this way)
      std::vector<unsigned char> bytes;
      FILE* f = fopen(scenePath.c_str(), "rb");
      if (!f) { std::cout << "cannot open " << scenePath << std::endl; return 1; }
      unsigned char chunk[65536]; size_t got;
      while ((got = fread(chunk, 1, sizeof(chunk), f)) > 0) bytes.insert(bytes.end(), chunk, chunk + got);
      fclose(f);
      CHECKED_CALL(scene.setHostSceneData(bytes.data(), bytes.size(), err));
    } else {
      CHECKED_CALL(scene.load(scenePath.c_str(), err));
    }
    const SceneHeader* sceneHeader = SCENE_HEADER(scene.getHostSceneData());
    std::cout << "Scene info:\n   Bounds: Min: X=" << sceneHeader->modelsBoundingBox.bounds[0].x << " Y=" << sceneHeader->modelsBoundingBox.bounds[0].y
              << " Z=" << sceneHeader->modelsBoundingBox.bounds[0].z << "\n           Max: X=" << sceneHeader->modelsBoundingBox.bounds[1].x
              << " Y=" << sceneHeader->modelsBoundingBox.bounds[1].y << " Z=" << sceneHeader->modelsBoundingBox.bounds[1].z
              << "\n   Num of Materials: " << sceneHeader->numberOfMaterials << "\n   Num of Lights: " << sceneHeader->numberOfLights
              << "\n   Num of Tris: " << sceneHeader->totalNumberOfTriangles << "\n   Total scene data size: " << sceneHeader->totalDataSize << " bytes" << std::endl;
    CHECKED_CALL(scene.loadToGPU(context, err));

    // camera (OpenCLRayTracerDemo.cpp:202-217)
    Camera camera;
    camera.resX = window_width; camera.resY = window_height;
    camera.FOVDistance = FOVDistFromAngle(90, (float)window_width, (float)window_height);
    camera.supersampilngFactor = 1;
    CL_FLOAT3 half, cameraPosition;
    half.x = (sceneHeader->modelsBoundingBox.bounds[1].x - sceneHeader->modelsBoundingBox.bounds[0].x) * 0.5f;
    half.y = (sceneHeader->modelsBoundingBox.bounds[1].y - sceneHeader->modelsBoundingBox.bounds[0].y) * 0.5f;
    half.z = (sceneHeader->modelsBoundingBox.bounds[1].z - sceneHeader->modelsBoundingBox.bounds[0].z) * 0.5f;
    cameraPosition.x = sceneHeader->modelsBoundingBox.bounds[0].x + half.x;
    cameraPosition.y = sceneHeader->modelsBoundingBox.bounds[0].y + half.y;
    cameraPosition.z = sceneHeader->modelsBoundingBox.bounds[0].z - half.z;
    cameraPosition.w = 0.0f;
    camera.viewTransform = identityTransform();
    Quaternion cameraOrientation = zeroRotation();

    std::shared_ptr<AccelerationStructureManager> accelerationStruct;
    if (accelerationStructInUse == BVH) accelerationStruct.reset(new BVHManager(context, scene));
    else if (accelerationStructInUse == GRID) accelerationStruct.reset(new TwoLevelGridManager(context, scene));
    else accelerationStruct.reset(new KDTreeManager(context, scene));
    CHECKED_CALL(accelerationStruct->initialize(err));
    CHECKED_CALL(accelerationStruct->initializeFrame(err));
    CHECKED_CALL(accelerationStruct->construct(err));
    std::cout << "Acceleration structure constructed" << std::endl;

    CLBuffer colors(context, pixelCount * 4, CLBufferFlags::ReadWrite);
    CLBuffer shadowRays(context, 16, CLBufferFlags::ReadWrite), shadowContacts(context, 16, CLBufferFlags::ReadWrite);
    std::vector<unsigned char> host(pixelCount * 4);
    const size_t frames = script.size() + 1;
    for (size_t f = 0; f < frames; f++) {
      if (f > 0) {                                         // appKeyboard / appMotion (:378-451)
        const char key = script[f - 1];
        const float delta = 0.3f, rot = 5.0f * (3.14159265359f / 180.0f);
        const Matrix4& t = camera.viewTransform;
        const float fw[3] = {t.data[2], t.data[6], t.data[10]}, sd[3] = {t.data[0], t.data[4], t.data[8]};
        CL_FLOAT3 r; r.x = r.y = r.z = r.w = 0.0f;
        switch (key) {
          case 'U': cameraPosition.x += fw[0] * delta; cameraPosition.y += fw[1] * delta; cameraPosition.z += fw[2] * delta; break;
          case 'D': cameraPosition.x -= fw[0] * delta; cameraPosition.y -= fw[1] * delta; cameraPosition.z -= fw[2] * delta; break;
          case 'L': cameraPosition.x += sd[0] * delta; cameraPosition.y += sd[1] * delta; cameraPosition.z += sd[2] * delta; break;
          case 'R': cameraPosition.x -= sd[0] * delta; cameraPosition.y -= sd[1] * delta; cameraPosition.z -= sd[2] * delta; break;
          case 'a': r.y = rot; rotateByVector(&cameraOrientation, r); break;
          case 'd': r.y = -rot; rotateByVector(&cameraOrientation, r); break;
          case 'w': r.x = rot; rotateByVector(&cameraOrientation, r); break;
          case 's': r.x = -rot; rotateByVector(&cameraOrientation, r); break;
          default: break;
        }
      }
      // RTFrame (:271-326)
      normalizeQuaternion(&cameraOrientation);
      setOrientationAndPos(&camera.viewTransform, cameraOrientation, cameraPosition);
      CHECKED_CALL(accelerationStruct->generateContacts(camera, err));
      cl_mem contacts = accelerationStruct->getPrimaryContacts()->getCLMem();
      cl_mem shadowC = NULL;
      if (shadows) {
        uint32_t perContact = 0;
        const size_t nl = (size_t)sceneHeader->numberOfLights + 1;
        shadowRays.resize(pixelCount * nl * sizeof(Ray)); shadowContacts.resize(pixelCount * nl * sizeof(Contact));
        if (clrt_generate_shadow_rays(context.handle(), scene.deviceHandle(), &camera, contacts, (uint32_t)pixelCount, shadowRays.getCLMem(), &perContact)) {
          std::cout << clrt_last_error() << std::endl; return 1;
        }
        CHECKED_CALL(accelerationStruct->generateContacts(shadowRays, shadowContacts, (unsigned)(pixelCount * perContact), err));
        shadowC = shadowContacts.getCLMem();
      }
      if (clrt_shade(context.handle(), scene.deviceHandle(), &camera, contacts, (uint32_t)pixelCount, shadowC, colors.getCLMem(), NULL)) {
        std::cout << clrt_last_error() << std::endl; return 1;
      }
      CHECKED_CALL(colors.copyToHost(host.data(), err));
      size_t lit = 0;
      for (size_t i = 0; i < pixelCount; i++) lit += host[4 * i + 3] != 0;
      std::cout << "frame " << f << ": " << lit << " of " << pixelCount << " pixels hit" << std::endl;
      if (!outPrefix.empty()) {
        char name[32]; snprintf(name, sizeof(name), "_%04zu.ppm", f);
        if (!write_ppm(outPrefix + name, host, window_width, window_height)) { std::cout << "cannot write " << outPrefix << name << std::endl; return 1; }
      }
    }
  } catch (const CLInterfaceException& e) {
    std::cout << static_cast<const Errata&>(e);
    return 1;
  }
  return 0;
}
