/* Minimal stand-in for the Khronos <CL/cl.h> host types, used ONLY to compile the
 * reference's dual-target headers (R/Include/CLData/**) on the host for the oracle
 * (oracle/ref_bridge.cpp).  TEST INFRASTRUCTURE - never linked into the product.
 *
 * Deliberate choices:
 *  - vector types are plain 4-byte-aligned PODs, so that reads from the (unpadded,
 *    mis-aligned) v0.6 scene buffer are defined behaviour on the host;
 *  - cl_float3 == cl_float4 and cl_uint3 == cl_uint4 exactly as in the Khronos header
 *    (3-vectors occupy 16 bytes).
 */
#ifndef CLRT_ORACLE_SHIM_CL_H
#define CLRT_ORACLE_SHIM_CL_H
#include <stdint.h>
typedef int16_t  cl_short;  typedef uint16_t cl_ushort;
typedef int32_t  cl_int;    typedef uint32_t cl_uint;
typedef int64_t  cl_long;   typedef uint64_t cl_ulong;
typedef float    cl_float;
#define CLRT_VEC(T, N2, N4, N3) \
  typedef struct { T x, y; } N2; \
  typedef struct { T x, y, z, w; } N4; \
  typedef N4 N3;
CLRT_VEC(cl_short,  cl_short2,  cl_short4,  cl_short3)
CLRT_VEC(cl_ushort, cl_ushort2, cl_ushort4, cl_ushort3)
CLRT_VEC(cl_int,    cl_int2,    cl_int4,    cl_int3)
CLRT_VEC(cl_uint,   cl_uint2,   cl_uint4,   cl_uint3)
CLRT_VEC(cl_long,   cl_long2,   cl_long4,   cl_long3)
CLRT_VEC(cl_ulong,  cl_ulong2,  cl_ulong4,  cl_ulong3)
CLRT_VEC(cl_float,  cl_float2,  cl_float4,  cl_float3)
#undef CLRT_VEC
#endif
