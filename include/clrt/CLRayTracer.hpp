// CLRayTracer.hpp — the C++ host side above the C ABI (include/clrt.h): the reference's own class names, method
// signatures, argument meaning and error convention, so that code written against kotturtech/OpenCLRayTracer
// (R = /root/reference/OpenCLRayTracer) compiles and runs against libclrt_cuda.so with its call sites unchanged.
//
//   CLRayTracer::Common::Result / Errata / CLInterfaceException      R/Include/Common/Errata.h:40-131
//   CLRayTracer::OpenCLUtils::CLExecutionContext                     R/Include/OpenCLUtils/CLExecutionContext.h (queue ops used by the managers)
//   CLRayTracer::OpenCLUtils::CLBuffer                               R/Include/OpenCLUtils/CLBuffer.h:39-112
//   CLRayTracer::Common::BitonicSort                                 R/Include/Algorithms/Sorting.h:54-85
//   CLRayTracer::Common::PrefixSum                                   R/Include/Algorithms/PrefixSum.h:48-83
//   CLRayTracer::Scene                                               R/Include/Scene/Scene.h:66-106   (load(): host/Scene.cpp -> libclrt_host)
//   CLRayTracer::AccelerationStructures::AccelerationStructureManager / BVHManager / TwoLevelGridManager
//                                                                    R/Include/Algorithms/AccelerationStructureManager.h:52-78,
//                                                                    BVHManager.h:59-126, TwoLevelGridManager.h:61-178
//   Camera, Contact, Ray (layouts of include/clrt_types.h), FOVDistFromAngle, quaternion camera helpers
//                                                                    R/Include/CLData/CLStructs.h:44-118, Transform.h:76-147,397-413
//
// Differences a caller can see: `cl_mem` is a device pointer; boost::shared_ptr is std::shared_ptr;
// constructors throw CLInterfaceException BY VALUE (the reference throws a heap pointer); calls return
// after the device finished, like the reference's enqueue -> clFlush -> clWaitForEvents (BVHManager.cpp:184-192).
// Header-only except Scene::load.  No OpenCL, no boost.  There is no CPU fallback: without a CUDA device the context
// constructor throws.
#ifndef CLRT_CLRAYTRACER_HPP
#define CLRT_CLRAYTRACER_HPP

#include <cfloat>
#include <cmath>
#include <cstring>
#include <exception>
#include <memory>
#include <ostream>
#include <sstream>
#include <string>

#include "../clrt.h"

typedef void* cl_mem;            // what a cl_mem becomes: a device pointer
typedef uint32_t cl_uint;
typedef uint64_t cl_ulong;

namespace CLRayTracer {

// ---- Common ------------------------------------------------------------------------------------------------------
namespace Common {

enum Result { Success = 0, Error = 1 };                                   // Errata.h:47-51

class Errata {                                                            // Errata.h:57-110
 public:
  Errata(const std::string& errorMessage, const std::string& file, const std::string& function, int line, const std::exception* ex = NULL)
      : _message(errorMessage), _file(file), _function(function), _line(line), _exception(ex) {}
  Errata() : _line(0), _exception(NULL) {}
  std::string message() const { return _message; }
  std::string file() const { return _file; }
  std::string function() const { return _function; }
  int line() const { return _line; }
  const std::exception* exception() const { return _exception; }

 protected:
  std::string _message, _file, _function;
  int _line;
  const std::exception* _exception;
};
inline std::ostream& operator<<(std::ostream& o, const Errata& err) {
  o << "Error: " << err.message() << "\n  in " << err.function() << " (" << err.file() << ":" << err.line() << ")\n";
  if (err.exception()) o << "  caused by: " << err.exception()->what() << "\n";
  return o;
}

class CLInterfaceException : public Errata, public std::exception {       // Errata.h:113-131
 public:
  CLInterfaceException(const std::string& errorMessage, const std::string& file, const std::string& function, int line, const std::exception* ex = NULL)
      : Errata(errorMessage, file, function, line, ex) {}
  virtual ~CLInterfaceException() throw() {}
  virtual const char* what() const throw() { return _message.c_str(); }
};

}  // namespace Common

#define CLRT_FILL_ERRATA(err, msg)                                                                           \
  {                                                                                                          \
    std::stringstream clrt_ss__;                                                                             \
    clrt_ss__ << msg;                                                                                        \
    (err) = CLRayTracer::Common::Errata(clrt_ss__.str(), std::string(__FILE__), std::string(__FUNCTION__), __LINE__); \
  }
#define CLRT_THROW(msg)                                                                                      \
  {                                                                                                          \
    std::stringstream clrt_ss__;                                                                             \
    clrt_ss__ << msg;                                                                                        \
    throw CLRayTracer::Common::CLInterfaceException(clrt_ss__.str(), std::string(__FILE__), std::string(__FUNCTION__), __LINE__); \
  }
// status of a C-ABI call -> Result + Errata (the message is clrt_last_error())
#define CLRT_CALL(expr, err)                               \
  if ((expr) != 0) {                                       \
    CLRT_FILL_ERRATA(err, #expr << ": " << clrt_last_error()); \
    return CLRayTracer::Common::Error;                     \
  }

// ---- CLData: layouts that cross the boundary (include/clrt_types.h) and the camera helpers ---------------------------
typedef clrt_camera Camera;          // CLStructs.h:44-51
typedef clrt_contact Contact;        // CLStructs.h:112-118 (pad[0] additionally carries the hit triangle id)
typedef clrt_ray Ray;                // CLStructs.h:82-87
typedef clrt_matrix4 Matrix4;
typedef clrt_float4 CL_FLOAT3;
typedef float CL_FLOAT;
struct CL_UINT3 { uint32_t x, y, z, w; };                                  // cl_uint3 occupies 16 bytes
typedef clrt_aabb AABB;                                                    // Primitives/AABB.h:47-52
typedef clrt_scene_header SceneHeader;
struct Quaternion { float r, i, j, k; };                                  // Transform.h:60-70

inline float FOVDistFromAngle(float angle, float /*resx*/, float resy) {  // CLStructs.h:70-74
  const float PI = 3.14159265359f, DEG2RAD = PI / 180.0f;
  const float a = resy * 0.5f;
  return a / (float)tan(angle * 0.5f * DEG2RAD);
}
inline Quaternion zeroRotation() { Quaternion q; q.r = 1.0f; q.i = q.j = q.k = 0.0f; return q; }
inline void normalizeQuaternion(Quaternion* q) {                          // Transform.h:88-106
  float d = q->r * q->r + q->i * q->i + q->j * q->j + q->k * q->k;
  if (d < FLT_EPSILON) { q->i = q->j = q->k = 0.0f; q->r = 1.0f; return; }
  d = 1.0f / (float)sqrt(d);
  q->r *= d; q->i *= d; q->j *= d; q->k *= d;
}
inline Quaternion mult(const Quaternion& a, const Quaternion& b) {        // Hamilton product, Transform.h:114-126
  Quaternion q;
  q.r = a.r * b.r - a.i * b.i - a.j * b.j - a.k * b.k;
  q.i = a.r * b.i + a.i * b.r + a.j * b.k - a.k * b.j;
  q.j = a.r * b.j + a.j * b.r + a.k * b.i - a.i * b.k;
  q.k = a.r * b.k + a.k * b.r + a.i * b.j - a.j * b.i;
  return q;
}
inline void rotateByVector(Quaternion* o, CL_FLOAT3 rot) {               // Transform.h:134-147
  Quaternion q; q.r = 0; q.i = rot.x; q.j = rot.y; q.k = rot.z;
  q = mult(q, *o);
  o->r += q.r * 0.5f; o->i += q.i * 0.5f; o->j += q.j * 0.5f; o->k += q.k * 0.5f;
}
inline Matrix4 identityTransform() {
  Matrix4 m; std::memset(&m, 0, sizeof(m)); m.data[0] = m.data[5] = m.data[10] = 1.0f; return m;
}
inline void setOrientationAndPos(Matrix4* t, const Quaternion& q, CL_FLOAT3 pos) {     // Transform.h:397-413
  t->data[0] = 1 - (2 * q.j * q.j + 2 * q.k * q.k); t->data[1] = 2 * q.i * q.j + 2 * q.k * q.r; t->data[2] = 2 * q.i * q.k - 2 * q.j * q.r; t->data[3] = pos.x;
  t->data[4] = 2 * q.i * q.j - 2 * q.k * q.r; t->data[5] = 1 - (2 * q.i * q.i + 2 * q.k * q.k); t->data[6] = 2 * q.j * q.k + 2 * q.i * q.r; t->data[7] = pos.y;
  t->data[8] = 2 * q.i * q.k + 2 * q.j * q.r; t->data[9] = 2 * q.j * q.k - 2 * q.i * q.r; t->data[10] = 1 - (2 * q.i * q.i + 2 * q.j * q.j); t->data[11] = pos.z;
}
inline const SceneHeader* SCENE_HEADER(const char* sceneData) { return reinterpret_cast<const SceneHeader*>(sceneData); }

// ---- OpenCLUtils ---------------------------------------------------------------------------------------------------
namespace OpenCLUtils {

namespace CLBufferFlags { enum CLBufferAccess { ReadWrite = 0, ReadOnly = 1, WriteOnly = 2 }; }

// One device + one in-order queue (CLExecutionContext.cpp:264-293).  `device` is the CUDA ordinal; `stream` (optional)
// lends an existing cudaStream_t.
class CLExecutionContext {
 public:
  explicit CLExecutionContext(int device = 0, void* stream = NULL) : _ctx(NULL) {
    if (clrt_ctx_create(device, stream, &_ctx) != 0) CLRT_THROW("CLExecutionContext: " << clrt_last_error());
  }
  ~CLExecutionContext() { if (_ctx) clrt_ctx_destroy(_ctx); }
  Common::Result finishQueue(Common::Errata& err) const { CLRT_CALL(clrt_ctx_sync(_ctx), err); return Common::Success; }
  Common::Result flushQueue(Common::Errata&) const { return Common::Success; }       // a CUDA stream needs no flush
  // blocking, like the reference's (CLExecutionContext.cpp:386-470)
  Common::Result enqueueWriteBuffer(const void* src, cl_mem dst, size_t bytes, Common::Errata& err) const {
    CLRT_CALL(clrt_memcpy_htod(_ctx, dst, src, bytes), err); return Common::Success;
  }
  Common::Result enqueueReadBuffer(cl_mem src, void* dst, size_t bytes, Common::Errata& err) const {
    CLRT_CALL(clrt_memcpy_dtoh(_ctx, dst, src, bytes), err); return Common::Success;
  }
  // device getters the managers query (CLInterface.h:183,284,343)
  cl_ulong getLocalMemSize() const { return info(CLRT_DEV_LOCAL_MEM_SIZE); }
  size_t getMaxWorkGroupSize() const { return (size_t)info(CLRT_DEV_MAX_WORK_GROUP_SIZE); }
  cl_uint getMaxComputeUnits() const { return (cl_uint)info(CLRT_DEV_MAX_COMPUTE_UNITS); }
  cl_ulong getGlobalMemSize() const { return info(CLRT_DEV_GLOBAL_MEM_SIZE); }
  clrt_ctx* handle() const { return _ctx; }

 private:
  cl_ulong info(int what) const { uint64_t v = 0; clrt_ctx_device_info(_ctx, what, &v); return v; }
  CLExecutionContext(const CLExecutionContext&);
  CLExecutionContext& operator=(const CLExecutionContext&);
  clrt_ctx* _ctx;
};

class CLBuffer {                                                          // CLBuffer.h:39-112, CLBuffer.cpp:42-122
 public:
  CLBuffer(const CLExecutionContext& context, size_t size, CLBufferFlags::CLBufferAccess access)
      : _context(context), _access(access), _buf(NULL) {
    if (clrt_buf_alloc(context.handle(), size, NULL, &_buf) != 0) CLRT_THROW("CLBuffer: " << clrt_last_error());
  }
  CLBuffer(const CLExecutionContext& context, size_t size, void* source, CLBufferFlags::CLBufferAccess access)
      : _context(context), _access(access), _buf(NULL) {
    if (clrt_buf_alloc(context.handle(), size, source, &_buf) != 0) CLRT_THROW("CLBuffer: " << clrt_last_error());
  }
  virtual ~CLBuffer() { if (_buf) clrt_buf_free(_buf); }
  void resize(size_t newSize) {                                           // grow-only, contents discarded on growth (CLBuffer.cpp:70-94)
    if (clrt_buf_resize(_buf, newSize) != 0) CLRT_THROW("CLBuffer::resize: " << clrt_last_error());
  }
  size_t getSize() const { return clrt_buf_size(_buf); }
  size_t getActualSize() const { return clrt_buf_capacity(_buf); }
  CLBufferFlags::CLBufferAccess getAccess() const { return _access; }
  cl_mem getCLMem() const { return clrt_buf_dptr(_buf); }
  Common::Result copyToHost(void* target, Common::Errata& err) const { return copyToHost(target, 0, err); }
  Common::Result copyToHost(void* target, size_t offset, Common::Errata& err) const {
    CLRT_CALL(clrt_buf_download(_buf, target, offset, clrt_buf_size(_buf) - offset), err); return Common::Success;
  }
  Common::Result fill(const void* pattern, size_t patternSize, Common::Errata& err) {   // CLExecutionContext::enqueueFillBuffer
    CLRT_CALL(clrt_buf_fill(_buf, pattern, patternSize, clrt_buf_size(_buf)), err); return Common::Success;
  }
  clrt_buf* handle() const { return _buf; }

 private:
  CLBuffer(const CLBuffer&);
  CLBuffer& operator=(const CLBuffer&);
  const CLExecutionContext& _context;
  CLBufferFlags::CLBufferAccess _access;
  clrt_buf* _buf;
};

}  // namespace OpenCLUtils

// ---- building blocks ---------------------------------------------------------------------------------------------
namespace Common {

class BitonicSort {                                                       // Sorting.h:54-85, BitonicSort.cpp:118-189
 public:
  // useKeyValue == true sorts uint2 (key, value) pairs by key — the only mode the reference's call sites use.
  // compatTieOrder: reproduce the reference's bitonic-network permutation among equal keys (needs n = 2^k).
  BitonicSort(const OpenCLUtils::CLExecutionContext& context, bool useKeyValue = true, bool compatTieOrder = false)
      : _context(context), _useKeyValue(useKeyValue), _compat(compatTieOrder) {}
  Result initialize(Errata& err) {
    if (!_useKeyValue) { CLRT_FILL_ERRATA(err, "BitonicSort: only key-value (uint2) sorting is provided"); return Error; }
    return Success;
  }
  Result sort(cl_mem input, size_t numberOfItems, Errata& err) const {
    CLRT_CALL(clrt_sort_pairs_u32(_context.handle(), input, numberOfItems, _compat ? CLRT_SORT_BITONIC_COMPAT : CLRT_SORT_RADIX), err);
    CLRT_CALL(clrt_ctx_sync(_context.handle()), err);
    return Success;
  }

 private:
  const OpenCLUtils::CLExecutionContext& _context;
  bool _useKeyValue, _compat;
};

class PrefixSum {                                                         // PrefixSum.h:48-83, PrefixSum.cpp:94-133 (INCLUSIVE)
 public:
  explicit PrefixSum(const OpenCLUtils::CLExecutionContext& context) : _context(context) {}
  Result initialize(Errata&) { return Success; }
  Result computePrefixSum(cl_mem input, cl_mem output, size_t numberOfItems, Errata& err) const {
    CLRT_CALL(clrt_scan_inclusive_u32(_context.handle(), input, output, numberOfItems), err);
    CLRT_CALL(clrt_ctx_sync(_context.handle()), err);
    return Success;
  }

 private:
  const OpenCLUtils::CLExecutionContext& _context;
};

}  // namespace Common

// ---- Scene ---------------------------------------------------------------------------------------------------------
class Scene {                                                             // Scene.h:66-106
 public:
  Scene() : _hostSceneData(NULL), _sceneDataSize(0), _device(NULL) {}
  virtual ~Scene() { if (_device) clrt_scene_destroy(_device); delete[] _hostSceneData; }
  // Scene.cpp:231-444 — scene description file (LIGHT / MESH lines; OBJ + MTL models) -> packed v0.6 scene buffer.
  // Implemented in host/Scene.cpp (libclrt_host).
  Common::Result load(const char* filename, Common::Errata& err);
  // adopts an already packed buffer (e.g. a synthetic scene)
  Common::Result setHostSceneData(const void* packed, cl_ulong bytes, Common::Errata& err) {
    if (!packed || bytes < sizeof(SceneHeader)) { CLRT_FILL_ERRATA(err, "Scene: packed buffer too small"); return Common::Error; }
    delete[] _hostSceneData;
    _hostSceneData = new char[bytes]; std::memcpy(_hostSceneData, packed, bytes); _sceneDataSize = bytes;
    return Common::Success;
  }
  Common::Result loadToGPU(const OpenCLUtils::CLExecutionContext& context, Common::Errata& err) {     // Scene.cpp:451-456
    if (!_hostSceneData) { CLRT_FILL_ERRATA(err, "Scene::loadToGPU: no scene loaded"); return Common::Error; }
    if (_device) { clrt_scene_destroy(_device); _device = NULL; }
    CLRT_CALL(clrt_scene_upload(context.handle(), _hostSceneData, _sceneDataSize, &_device), err);
    return Common::Success;
  }
  cl_mem getDeviceSceneData() const { return _device ? clrt_scene_dptr(_device) : NULL; }
  const char* getHostSceneData() const { return _hostSceneData; }
  cl_ulong getSceneDataSize() const { return _sceneDataSize; }
  clrt_scene* deviceHandle() const { return _device; }

 private:
  Scene(const Scene&);
  Scene& operator=(const Scene&);
  char* _hostSceneData;
  cl_ulong _sceneDataSize;
  clrt_scene* _device;
};

// ---- acceleration structures -----------------------------------------------------------------------------------------
namespace AccelerationStructures {

class AccelerationStructureManager {                                      // AccelerationStructureManager.h:52-78
 public:
  AccelerationStructureManager(const OpenCLUtils::CLExecutionContext& context, const Scene& scene)
      : _context(context), _scene(scene), _accel(NULL) {}
  virtual Common::Result initialize(Common::Errata& err) = 0;
  virtual Common::Result initializeFrame(Common::Errata& err) {
    if (!_accel) { CLRT_FILL_ERRATA(err, "initializeFrame: initialize() was not called"); return Common::Error; }
    CLRT_CALL(clrt_accel_init_frame(_accel), err);
    return Common::Success;
  }
  virtual Common::Result construct(Common::Errata& err) {
    if (!_accel) { CLRT_FILL_ERRATA(err, "construct: initialize() was not called"); return Common::Error; }
    CLRT_CALL(clrt_accel_construct(_accel), err);
    CLRT_CALL(clrt_ctx_sync(_context.handle()), err);                     // the reference returns after clWaitForEvents
    return Common::Success;
  }
  virtual Common::Result generateContacts(Camera& cam, Common::Errata& err) {
    if (!_accel) { CLRT_FILL_ERRATA(err, "generateContacts: initialize() was not called"); return Common::Error; }
    // the contacts live in a CLBuffer the manager owns and resizes per frame (BVHManager.cpp:240-244)
    const size_t bytes = (size_t)cam.resX * cam.resY * sizeof(Contact);
    try {
      if (!_primaryContactsArray) _primaryContactsArray.reset(new OpenCLUtils::CLBuffer(_context, bytes ? bytes : sizeof(Contact), OpenCLUtils::CLBufferFlags::ReadWrite));
      else _primaryContactsArray->resize(bytes);
    } catch (const Common::CLInterfaceException& e) { err = e; return Common::Error; }
    CLRT_CALL(clrt_trace_primary(_accel, &cam, _primaryContactsArray->getCLMem()), err);
    CLRT_CALL(clrt_accel_check(_accel), err);              // waits for the queue; fails on a traversal-stack overflow
    return Common::Success;
  }
  virtual Common::Result generateContacts(OpenCLUtils::CLBuffer& rays, OpenCLUtils::CLBuffer& contacts, const unsigned int rayCount, Common::Errata& err) {
    if (!_accel) { CLRT_FILL_ERRATA(err, "generateContacts: initialize() was not called"); return Common::Error; }
    if (rays.getSize() < (size_t)rayCount * sizeof(Ray) || contacts.getSize() < (size_t)rayCount * sizeof(Contact)) {
      CLRT_FILL_ERRATA(err, "generateContacts: buffers are smaller than rayCount elements"); return Common::Error;
    }
    CLRT_CALL(clrt_trace_rays(_accel, rays.getCLMem(), contacts.getCLMem(), rayCount), err);
    CLRT_CALL(clrt_accel_check(_accel), err);
    return Common::Success;
  }
  // AccelerationStructureManager.h:75, BVHManager.h:108, TwoLevelGridManager.h:120
  virtual const std::shared_ptr<OpenCLUtils::CLBuffer> getPrimaryContacts() const { return _primaryContactsArray; }
  virtual ~AccelerationStructureManager() { if (_accel) clrt_accel_destroy(_accel); }

  // additions (not in the reference): traversal policy, trace tree policy, stage timings
  Common::Result setTraceMode(int mode, Common::Errata& err) { CLRT_CALL(clrt_accel_set_trace_mode(_accel, mode), err); return Common::Success; }
  clrt_accel* handle() const { return _accel; }

 protected:
  const OpenCLUtils::CLExecutionContext& _context;
  const Scene& _scene;
  clrt_accel* _accel;
  std::shared_ptr<OpenCLUtils::CLBuffer> _primaryContactsArray;
};

class BVHManager : public AccelerationStructureManager {                   // BVHManager.h:59-126
 public:
  BVHManager(const OpenCLUtils::CLExecutionContext& context, const Scene& scene) : AccelerationStructureManager(context, scene) {}
  virtual Common::Result initialize(Common::Errata& err) {                // BVHManager.cpp:74-116 (kernels are compiled ahead of time here)
    if (!_scene.deviceHandle()) { CLRT_FILL_ERRATA(err, "BVHManager::initialize: Scene::loadToGPU was not called"); return Common::Error; }
    if (_accel) { clrt_accel_destroy(_accel); _accel = NULL; }
    CLRT_CALL(clrt_bvh_create(_context.handle(), _scene.deviceHandle(), &_accel), err);
    return Common::Success;
  }
};

class TwoLevelGridManager : public AccelerationStructureManager {          // TwoLevelGridManager.h:61-178
 public:
  TwoLevelGridManager(const OpenCLUtils::CLExecutionContext& context, const Scene& scene)
      : AccelerationStructureManager(context, scene), _topLevelDensity(2.0f), _leafDensity(2.0f) {}
  virtual Common::Result initialize(Common::Errata& err) {
    if (!_scene.deviceHandle()) { CLRT_FILL_ERRATA(err, "TwoLevelGridManager::initialize: Scene::loadToGPU was not called"); return Common::Error; }
    if (_accel) { clrt_accel_destroy(_accel); _accel = NULL; }
    CLRT_CALL(clrt_tlg_create(_context.handle(), _scene.deviceHandle(), _topLevelDensity, _leafDensity, &_accel), err);
    return Common::Success;
  }
  // Properties (TwoLevelGridManager.h:112-119): plain members; like the reference's, they are read when the grid data is
  // calculated, i.e. at the next initializeFrame() (TwoLevelGridManager.cpp:160-201 -> calculateGridData :584-600).
  void setTopLevelDensity(CL_FLOAT value) { _topLevelDensity = value; if (_accel) clrt_tlg_set_densities(_accel, _topLevelDensity, _leafDensity); }
  void setLeafDensity(CL_FLOAT value) { _leafDensity = value; if (_accel) clrt_tlg_set_densities(_accel, _topLevelDensity, _leafDensity); }
  CL_FLOAT getTopLevelDensity() const { return _topLevelDensity; }
  CL_FLOAT getLeafDensity() const { return _leafDensity; }
  // Utility functions (TwoLevelGridManager.h:129,134; .cpp:566-578,606-609)
  CL_UINT3 getResolution() const {
    CL_UINT3 r; r.x = r.y = r.z = r.w = 0;
    uint32_t v[3] = {0, 0, 0};
    if (_accel && clrt_tlg_resolution(_accel, v) == 0) { r.x = v[0]; r.y = v[1]; r.z = v[2]; }
    return r;
  }
  AABB getBounds() const { return SCENE_HEADER(_scene.getHostSceneData())->modelsBoundingBox; }
  // ours: the grid data of the last initializeFrame() as the kernels use it (resolution, steps, box, leaf density)
  Common::Result getGridData(clrt_grid_data& out, Common::Errata& err) const {
    CLRT_CALL(clrt_tlg_export(_accel, &out, NULL, NULL, NULL, NULL, NULL), err);
    return Common::Success;
  }

 private:
  float _topLevelDensity, _leafDensity;
};

// The reference's TODO 2.1 (KNOWN_ISSUES_AND_TODO.txt:24-26: "Adding KD trees - Construction may be based on method which is used to
// construct the BVH") as a third manager behind the plug-in interface (AccelerationStructureManager.h:52-78).
class KDTreeManager : public AccelerationStructureManager {
 public:
  KDTreeManager(const OpenCLUtils::CLExecutionContext& context, const Scene& scene) : AccelerationStructureManager(context, scene) {}
  virtual Common::Result initialize(Common::Errata& err) {
    if (!_scene.deviceHandle()) { CLRT_FILL_ERRATA(err, "KDTreeManager::initialize: Scene::loadToGPU was not called"); return Common::Error; }
    if (_accel) { clrt_accel_destroy(_accel); _accel = NULL; }
    CLRT_CALL(clrt_kdtree_create(_context.handle(), _scene.deviceHandle(), &_accel), err);
    return Common::Success;
  }
};

}  // namespace AccelerationStructures
}  // namespace CLRayTracer

#endif  // CLRT_CLRAYTRACER_HPP
