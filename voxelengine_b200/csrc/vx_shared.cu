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

extern "C" int vx_device_fill(void* d_dst, int byte, size_t bytes, int device, void* stream) {
    VX_REQUIRE(d_dst != nullptr, "NULL argument");
    DeviceGuard guard(device);
    VX_CUDA(cudaMemsetAsync(d_dst, byte, bytes, (cudaStream_t)stream));
    VX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return VX_OK;
}

extern "C" int vx_device_to_host(void* h_dst, const void* d_src, size_t bytes, int device, void* stream) {
    VX_REQUIRE(h_dst && d_src, "NULL argument");
    DeviceGuard guard(device);
    VX_CUDA(cudaMemcpyAsync(h_dst, d_src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
    VX_CUDA(cudaStreamSynchronize((cudaStream_t)stream));
    return VX_OK;
}

// ---- the frame in host memory shared by the ranks' processes ---------------------------------------------------------
// For a consumer on the host (main.cu:165-167 copies the frame down every frame) the bands need not meet on one GPU at
// all: every rank copies the bands it rendered straight into one host frame (POSIX shared memory, pinned in each
// process by vx_host_register), down its own PCIe link, all ranks at once.
extern "C" int vx_host_register(void* h_ptr, size_t bytes) {
    VX_REQUIRE(h_ptr && bytes, "NULL argument or empty buffer");
    VX_CUDA(cudaHostRegister(h_ptr, bytes, cudaHostRegisterPortable));
    return VX_OK;
}

extern "C" int vx_host_unregister(void* h_ptr) {
    VX_REQUIRE(h_ptr != nullptr, "NULL argument");
    VX_CUDA(cudaHostUnregister(h_ptr));
    return VX_OK;
}

extern "C" int vx_bands_to_host(void* h_frame, const void* d_frame, uint32_t W, uint32_t H, uint32_t band_count, uint32_t band_index,
                                int device, void* stream) {
    VX_REQUIRE(h_frame && d_frame, "NULL argument");
    VX_REQUIRE(W > 0 && H > 0, "empty frame");
    if (band_count == 0) { band_count = 1; band_index = 0; }
    VX_REQUIRE(band_index < band_count, "band_index must be below band_count");
    DeviceGuard guard(device);
    const size_t band_bytes = (size_t)W * 4u * 8u, row_bytes = (size_t)W * 4u;
    const uint32_t bands = (H + 7u) / 8u;
    if (band_index >= bands) return VX_OK;
    const uint32_t mine = (bands - band_index + band_count - 1u) / band_count;
    const uint32_t last = band_index + (mine - 1u) * band_count;                  // my last band may be cut by the frame's end
    const bool ragged = (size_t)(last + 1u) * 8u > H;
    const uint32_t full = ragged ? mine - 1u : mine;
    const size_t off = (size_t)band_index * band_bytes, pitch = (size_t)band_count * band_bytes;
    cudaStream_t cs = (cudaStream_t)stream;
    if (full) VX_CUDA(cudaMemcpy2DAsync((char*)h_frame + off, pitch, (const char*)d_frame + off, pitch, band_bytes, full, cudaMemcpyDeviceToHost, cs));
    if (ragged) {
        const size_t o = (size_t)last * band_bytes;
        VX_CUDA(cudaMemcpyAsync((char*)h_frame + o, (const char*)d_frame + o, ((size_t)H - (size_t)last * 8u) * row_bytes, cudaMemcpyDeviceToHost, cs));
    }
    return VX_OK;
}
