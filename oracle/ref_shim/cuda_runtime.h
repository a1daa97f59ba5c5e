/* Minimal stand-in for <cuda_runtime.h> so that the reference's cuda_noise.cuh compiles as plain
 * C++ with g++ (oracle/_ref/libvxref_noise.so).  Test infrastructure only. */
#ifndef VX_REF_SHIM_CUDA_RUNTIME_H
#define VX_REF_SHIM_CUDA_RUNTIME_H
#include <math.h>
#define __device__
#define __host__
#define __constant__
#define __forceinline__ inline
struct float3 { float x, y, z; };
static inline float3 make_float3(float x, float y, float z) { float3 v; v.x = x; v.y = y; v.z = z; return v; }
static inline float __saturatef(float v) { return v < 0.0f ? 0.0f : (v > 1.0f ? 1.0f : v); }
#endif
