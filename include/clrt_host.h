/* clrt_host.h — C entry points of libclrt_host.so, the host-side (no GPU) part of the drop-in: the scene loader/packer
 * behind CLRayTracer::Scene::load (R/src/CLRayTracerEngine/Scene.cpp:231-444, R = /root/reference/OpenCLRayTracer).
 * The C++ interface is include/clrt/CLRayTracer.hpp; these plain-C calls exist so that non-C++ hosts (the ctypes tests) can
 * reach the same code. */
#ifndef CLRT_HOST_H
#define CLRT_HOST_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif
/* Scene::load: scene description file (LIGHT / SPHERE / MESH lines; Wavefront OBJ + MTL models) -> packed v0.6 scene buffer
 * (format: R/Include/CLData/SceneBufferParser.h).  0 on success; *out_buf is released with clrt_host_free. */
int clrt_host_scene_load(const char* scene_file, void** out_buf, uint64_t* out_bytes);
void clrt_host_free(void* buf);
const char* clrt_host_last_error(void);
#ifdef __cplusplus
}
#endif
#endif
