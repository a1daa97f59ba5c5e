// vx_host.h — host-side objects behind the C ABI (include/vxrt.h) and error plumbing.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include "../../include/vxrt.h"
#include "vx_device.cuh"

namespace vx {

void set_error(const std::string& msg);
int cuda_fail(cudaError_t e, const char* what, const char* file, int line);

#define VX_CUDA(call)                                                                   \
    do {                                                                                \
        cudaError_t vx_e_ = (call);                                                     \
        if (vx_e_ != cudaSuccess) return vx::cuda_fail(vx_e_, #call, __FILE__, __LINE__); \
    } while (0)

#define VX_REQUIRE(cond, msg)                 \
    do {                                      \
        if (!(cond)) {                        \
            vx::set_error(msg);               \
            return VX_ERR_INVALID;            \
        }                                     \
    } while (0)

// Restores the caller's current device on scope exit.
struct DeviceGuard {
    int prev = -1;
    bool ok = false;
    explicit DeviceGuard(int device) {
        if (cudaGetDevice(&prev) != cudaSuccess) prev = -1;
        ok = cudaSetDevice(device) == cudaSuccess;
    }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

} // namespace vx

struct vx_world {
    // n / d == __umulhi(n, magic) >> shift for every n < 2^31 (d >= 2): magic = ceil(2^(31+L) / d), L = ceil(log2 d),
    // error term n * (magic*d - 2^(31+L)) < 2^31 * d <= 2^(31+L)
    static void magic_for(uint32_t d, uint32_t& magic, uint32_t& shift) {
        uint32_t L = 0;
        while ((1ull << L) < d) ++L;
        const unsigned __int128 num = (unsigned __int128)1 << (31 + L);
        magic = (uint32_t)((num + d - 1) / d);
        shift = L - 1;
    }
    int device = 0;
    uint32_t X = 0, Y = 0, Z = 0;
    int factor = 8;
    uint32_t cx = 0, cy = 0, cz = 0;
    uint64_t nbricks = 0, occupied = 0, solid_voxels = 0;
    uint32_t words_per_brick = 16;
    uint32_t* d_top = nullptr;   // (nbricks+31)/32 words
    uint2* d_records = nullptr;  // nbricks, tiled order (build-time only; freed once the padded copy exists)
    uint32_t* d_ptop = nullptr;  // padded linear occupancy bits (traversal)
    uint2* d_precords = nullptr; // padded linear records (traversal)
    uint32_t* d_pool = nullptr;  // pool_capacity * words_per_brick; slots >= pool_used are zero
    uint64_t pool_capacity = 0, pool_used = 0; // in bricks; vx_world_edit hands out slots pool_used.. and grows the pool
    float build_ms = 0.0f, tiles_ms = 0.0f;
    vx::WorldView view() const {
        vx::WorldView v;
        v.top = d_top; v.records = d_records; v.pool = d_pool;
        v.cx = (int)cx; v.cy = (int)cy; v.cz = (int)cz; v.factor = factor;
        v.inv_factor = (factor & (factor - 1)) == 0 ? 1.0f / (float)factor : 0.0f;
        v.words_per_brick = words_per_brick; v.tiles_x = cx / 8; v.tiles_xy = (cx / 8) * (cy / 8);
        v.ptop = d_ptop; v.precords = d_precords; v.pnx = (int)cx + 2; v.pnxy = ((int)cx + 2) * ((int)cy + 2);
        magic_for((uint32_t)v.pnx, v.magic_x, v.shift_x);
        magic_for((uint32_t)v.pnxy, v.magic_xy, v.shift_xy);
        return v;
    }
};

namespace vx { struct BatchPipe; void destroy_batch_pipe(BatchPipe*); }
constexpr unsigned kNextSlots = 64;
struct vx_tracer {
    vx_world* world = nullptr;
    int device = 0;
    size_t max_rays = 0;
    // staging for the host-buffer entry points
    float* d_origins = nullptr;
    float* d_dirs = nullptr;
    uint8_t* d_hit = nullptr;
    float* d_pos = nullptr;
    float* d_normal = nullptr;
    int32_t* d_steps = nullptr;
    int32_t* d_voxel = nullptr;
    uint32_t* d_brick = nullptr;
    int32_t* d_csteps = nullptr;
    int32_t* d_fsteps = nullptr;
    uint32_t* d_counters = nullptr;
    void* d_frame = nullptr; size_t frame_bytes = 0; bool frame_seeded = false; // vx_render_host
    float tan_arg = -1.0f, tan_val = 0.0f;           // cached device tanf
    cudaStream_t stream = nullptr;
    unsigned long long* d_next = nullptr;            // ring of ray counters for the persistent kernels
    unsigned next_seq = 0;
    void* d_queue_rec = nullptr; void* d_queue_aux = nullptr; void* d_queue_lanes = nullptr;   // shadow queue of vx_render (one frame in flight per tracer)
    unsigned long long queue_cap = 0; size_t queue_aux_words = 0;
    int sm_count = 0;
    cudaStream_t copy_stream = nullptr; cudaEvent_t slice_ev[4] = {nullptr, nullptr, nullptr, nullptr}; // vx_render_host: copy/render overlap
    int fused_shade = 0;  // vx_tracer_set_fused_shade
    bool primary_only = false; unsigned long long* last_queue_count = nullptr; // vx_bounce_rays reuses vx_render's primary pass
    cudaStream_t l2_stream = nullptr; const void* l2_base = nullptr; bool l2_tried = false; // pin_top_level (vx_trace.cu)
    uint32_t* d_tile_trace = nullptr; size_t tile_trace_words = 0, tile_trace_tiles = 0; // VX_TILE_TRACE diagnostics (vx_trace.cu)
    vx::BatchPipe* batch = nullptr; // vx_trace: pinned staging ring, copy threads and streams of the host-array batch call
    bool timing = false; int timed_kernels = 0; cudaEvent_t ev[5] = {nullptr, nullptr, nullptr, nullptr, nullptr}; // vx_tracer_set_timing
};
