/* clrt_types.h — POD layouts that cross the drop-in boundary.
 *
 * These are the byte layouts of the reference's dual-target structs (R = /root/reference/OpenCLRayTracer);
 * sizes are checked against the reference's own headers compiled on the host (tests/test_oracle_pin.py).
 * Every 3-vector occupies 16 bytes (OpenCL float3 == float4).
 *
 *   clrt_bvh_node   R/Include/CLData/AccelerationStructs/BVHData.h:70-74   48 B
 *   clrt_aabb       R/Include/CLData/Primitives/AABB.h:47-52               32 B
 *   clrt_contact    R/Include/CLData/CLStructs.h:112-118                   32 B
 *   clrt_ray        R/Include/CLData/CLStructs.h:82-87                     48 B
 *   clrt_camera     R/Include/CLData/CLStructs.h:44-51                     64 B
 *   clrt_matrix4    R/Include/CLData/Transform.h (struct Matrix4)          48 B
 *   clrt_grid_data / clrt_top_cell  R/Include/CLData/AccelerationStructs/TwoLevelGridData.h:39-61   64 B / 16 B
 *   clrt_scene_header / clrt_model_header / clrt_mesh_header  R/Include/CLData/SceneBufferParser.h:59-70,133-153
 *   clrt_material   R/Include/CLData/Primitives/Material.h:46-58           96 B
 *   clrt_light      R/Include/CLData/Primitives/Light.h                    16 B
 */
#ifndef CLRT_TYPES_H
#define CLRT_TYPES_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct { float x, y, z, w; } clrt_float4;   /* also float3 (w = padding) */
typedef struct { uint32_t x, y; } clrt_uint2;       /* (key, value) pair of the sort building block */

typedef struct { clrt_float4 bounds[2]; } clrt_aabb;            /* [0] = min, [1] = max */

/* data[0] = parent; leaf: data[1..3] = triangle-in-submesh, submesh, model; inner: data[1..2] = childA, childB.
 * node type lives in bounds[0].w as a float (1.0f leaf, 0.0f inner) — BVHData.h:41-65. */
typedef struct { uint32_t data[4]; clrt_aabb boundingBox; } clrt_bvh_node;

/* pad[0] carries the hit primitive (global triangle) id — additive, layout preserving.
 * pad[1] is zero. */
typedef struct { uint32_t pixelIndex, materialIndex, pad[2]; clrt_float4 normalAndintersectionDistance; } clrt_contact;

typedef struct { uint32_t idx, _pad[3]; clrt_float4 origin, direction; } clrt_ray;

typedef struct { float data[12]; } clrt_matrix4;   /* 3x4 row major, translation in data[3], [7], [11] */
typedef struct { float FOVDistance; uint32_t resX, resY, supersampilngFactor; clrt_matrix4 viewTransform; } clrt_camera;

typedef struct { clrt_float4 posAndEnergy; } clrt_light;
typedef struct {
  clrt_float4 ambient, diffuse, specular, transmittance, emission;
  float shininess, ior, dissolve, illum;
} clrt_material;

typedef struct {
  uint64_t totalDataSize, numberOfPrimitives, numberOfLights, numberOfSpheres, numberOfMaterials,
           modelBufferSize, numberOfModels, totalNumberOfTriangles;
  clrt_aabb modelsBoundingBox;
} clrt_scene_header;                                                /* 96 B */
typedef struct { uint64_t dataSize, numberOfSubmeshes, numberOfTriangles, pad; clrt_aabb boundingBox; } clrt_model_header; /* 64 B */
typedef struct { uint64_t dataSize, numberOfTriangles, numberOfVertices, numberOfIndices, materialIndex, pad; } clrt_mesh_header; /* 48 B */

/* Two-Level Grid (TwoLevelGridData.h:39-61) */
typedef struct {
  uint32_t resX, resY, resZ;        /* top-level resolution */
  float stepX, stepY, stepZ;        /* top-level cell extent per axis */
  float leafDensity, padTo64;
  clrt_aabb box;                    /* scene box */
} clrt_grid_data;                   /* 64 B */
typedef struct { uint32_t resX, resY, resZ, firstLeafIdx; } clrt_top_cell;   /* 16 B */

/* Device-side flattened triangle produced at scene upload (ours; replaces the per-fetch header walks,
 * SceneBufferParser.h:175-195,253-293).  w lanes carry ids as raw uint32 bits. */
typedef struct {
  clrt_float4 v0;   /* w = materialIndex (MeshHeader.materialIndex) */
  clrt_float4 v1;   /* w = submesh index inside its model           */
  clrt_float4 v2;   /* w = model index                              */
} clrt_triangle;    /* 48 B */

#ifdef __cplusplus
}
#endif
#endif
