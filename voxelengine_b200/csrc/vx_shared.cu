// Shared frame for the multi-GPU partition: rank 0 owns the frame buffer, every other rank's process maps
// it over CUDA IPC and its frame kernels store their pixels straight into rank 0's HBM over NVLink / NVSwitch peer
// memory.  The reference has no multi-GPU path (one cudaMalloc'd framebuffer, main.cu:95-99); this is the exchange
// step of our frame partition, fused into the kernels' own pixel stores: there is no gather kernel and no staging copy.
#include "vx_host.h"

#include <cstring>

using vx::DeviceGuard;

static_assert(sizeof(cudaIpcMemHandle_t) == VX_SHARED_HANDLE_BYTES, "vx_shared handle size");

extern "C" int vx_shared_alloc(size_t bytes, int device, void** d_ptr, uint8_t handle[VX_SHARED_HANDLE_BYTES]) {
    VX_REQUIRE(d_ptr && handle && bytes, "NULL argument or empty buffer");
    DeviceGuard guard(device);
    void* p = nullptr;
    VX_CUDA(cudaMalloc(&p, bytes)); // its own allocation: the IPC handle names the allocation base
    cudaError_t e = cudaMemset(p, 0, bytes);
    cudaIpcMemHandle_t h;
    if (e == cudaSuccess) e = cudaIpcGetMemHandle(&h, p);
    if (e != cudaSuccess) { cudaFree(p); VX_CUDA(e); }
    std::memcpy(handle, &h, sizeof h);
    *d_ptr = p;
    return VX_OK;
}

extern "C" int vx_shared_open(const uint8_t handle[VX_SHARED_HANDLE_BYTES], int device, void** d_ptr) {
    VX_REQUIRE(d_ptr && handle, "NULL argument");
    DeviceGuard guard(device);
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void* p = nullptr;
    VX_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
    *d_ptr = p;
    return VX_OK;
}

extern "C" int vx_shared_close(void* d_ptr, int device) {
    if (!d_ptr) return VX_OK;
    DeviceGuard guard(device);
    VX_CUDA(cudaIpcCloseMemHandle(d_ptr));
    return VX_OK;
}

extern "C" int vx_shared_free(void* d_ptr, int device) {
    if (!d_ptr) return VX_OK;
    DeviceGuard guard(device);
    VX_CUDA(cudaFree(d_ptr));
    return VX_OK;
}

extern "C" int vx_device_to_host(void* h_dst, const void* d_src, size_t bytes, int device, void* stream) {
    VX_REQUIRE(h_dst && d_src, "NULL argument");
    DeviceGuard guard(device);
    VX_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    VX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return VX_OK;
}
