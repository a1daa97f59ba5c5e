/* Force-included (nvcc -include) before the reference's VolumeRaytracer.cu when building
 * oracle/_ref/libvxref_host.so: every `__device__` function of the reference becomes
 * `__host__ __device__`, so the reference's own march runs on the CPU unmodified.
 * Test infrastructure only. */
#include <cuda_runtime.h>
#undef __device__
#define __device__ __location__(host) __location__(device)
