// VoxelRT.h — header-compatible C++ shim: the reference's public names for the hot path
// (GPUDDA::VoxelRaytracer3D, GPUDDA::Graphics::*, CreateVoxels, GenerateLowresVoxelBuffer) implemented on
// the C ABI of libvxrt.so (include/vxrt.h).  VoxelApp/main.cu minus its SDL calls compiles against the three
// wrapper headers next to this file (VolumeRaytracer.cuh, Renderer.cuh, VoxelWorldBuilder.cuh) unchanged; see
// examples/headless_main.cu and INTEGRATION.md.
//
// Differences a caller can observe (all documented in INTEGRATION.md):
//  * CreateVoxels does not materialise the dense host grid (the world is built on the device in
//    GenerateLowresVoxelBuffer); `delete[] buffer.grid.Raw()` stays valid.
//  * The per-brick arrays returned by GenerateLowresVoxelBuffer are opaque placeholders; the brickmap lives on
//    the GPU behind the VoxelBuffer3D handle.
//  * The device pointers returned by GetVoxelBuffer*/GetVoxelBufferDataBounds are opaque.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <limits>
#include <memory>
#include <tuple>
#include <vector>
#include "../vxrt.h"

constexpr auto FLT_INF = std::numeric_limits<float>::infinity();

namespace GPUDDA {

inline void vxrt_check(int rc, const char* what) {
    if (rc != VX_OK) { // reference behaviour on CUDA errors: print and exit (Renderer.cuh:15-23)
        fprintf(stderr, "vxrt error %d in %s: %s\n", rc, what, vx_last_error());
        exit(EXIT_FAILURE);
    }
}

// VolumeRaytracer.cuh:107-137 (tiled-linear mode), 64-bit safe
inline uint64_t GetSampleIndex(uint32_t x, uint32_t y, uint32_t z, uint32_t width, uint32_t height) {
    const uint64_t tW = width / 8, tH = height / 8;
    return ((x / 8) + (y / 8) * tW + (uint64_t)(z / 8) * tW * tH) * 512 + (x % 8) + (y % 8) * 8 + (z % 8) * 64;
}

struct BitArray { // VolumeRaytracer.cuh:210-223 (host side only)
    size_t size = 0;
    uint32_t* data = nullptr;
    BitArray() = default;
    BitArray(size_t num_bits, bool /*isGPU*/) : size(num_bits), data(new uint32_t[(num_bits + 31) / 32]()) {}
    bool operator[](size_t i) const { return i < size && ((data[i / 32] >> (i % 32)) & 1u); }
    void set(size_t i, bool v) { if (v) data[i / 32] |= 1u << (i % 32); else data[i / 32] &= ~(1u << (i % 32)); }
    uint32_t* Raw() { return data; }
    size_t BitSize() const { return size; }
    size_t ByteSize() const { return (size + 31) / 32 * sizeof(uint32_t); }
};

template <size_t D> struct VoxelBuffer {
    BitArray grid{};
    uint16_t dimensions[D]{};
    // shim state
    vx_world* world = nullptr;     // set on the coarse buffer returned by GenerateLowresVoxelBuffer
    bool from_noise = false;       // set by CreateVoxels: the dense grid was never materialised
    uint32_t full_dims[3] = {0, 0, 0};
};
typedef VoxelBuffer<3> VoxelBuffer3D;
template <class T> struct Bounds { T min, max; };
typedef Bounds<float3> Bounds3Df;
constexpr size_t MAX_STEPS = 2048;

template <typename T> class RayTraceResults { // VolumeRaytracer.cuh:179-202
public:
    std::shared_ptr<bool[]> valid{};
    std::shared_ptr<T[]> hitPoint{};
    std::shared_ptr<T[]> normal{};
    std::shared_ptr<float[]> distance{};
    std::shared_ptr<int[]> voxelIndex{};
    std::shared_ptr<int[]> steps{};
    // parity planes the reference keeps as locals (VolumeRaytracer.cu:419, HitCell arithmetic)
    std::shared_ptr<int[]> voxel{};       // 3 per ray
    std::shared_ptr<uint32_t[]> brick{};
    RayTraceResults(size_t count) {
        if (!count) return;
        valid.reset(new bool[count]()); hitPoint.reset(new T[count]()); normal.reset(new T[count]());
        distance.reset(new float[count]()); voxelIndex.reset(new int[count]()); steps.reset(new int[count]());
        voxel.reset(new int[3 * count]()); brick.reset(new uint32_t[count]());
    }
};

class VoxelRaytracer3D { // VolumeRaytracer.cuh:291-377
    VoxelRaytracer3D(const VoxelRaytracer3D&) = delete;
    VoxelRaytracer3D& operator=(const VoxelRaytracer3D&) = delete;
    RayTraceResults<float3> resultsCPU = RayTraceResults<float3>(0);
    size_t count_ = 0;
    int factor = 1;
    vx_world* world_ = nullptr;
    vx_tracer* tracer_ = nullptr;
    std::vector<uint8_t> hit_;
public:
    explicit VoxelRaytracer3D(size_t count) : count_(count) { resultsCPU = RayTraceResults<float3>(count); hit_.resize(count); }
    ~VoxelRaytracer3D() { Free(); }
    void UploadVoxelBuffer(const VoxelBuffer3D& buff) { // attaches the device brickmap
        if (!buff.world) { fprintf(stderr, "UploadVoxelBuffer: buffer does not come from GenerateLowresVoxelBuffer\n"); exit(EXIT_FAILURE); }
        world_ = buff.world;
        if (tracer_) vx_tracer_destroy(tracer_);
        vxrt_check(vx_tracer_create(&tracer_, world_, count_), "vx_tracer_create");
    }
    void UploadVoxelBufferDatas(VoxelBuffer3D*, size_t) {}      // brick masks already live in the pool
    void UploadVoxelBufferDataBounds(Bounds3Df*, size_t) {}     // bounds already live in the records
    VoxelBuffer3D* GetVoxelBuffer() { return reinterpret_cast<VoxelBuffer3D*>(world_); }
    VoxelBuffer3D* GetVoxelBufferDatas() { return reinterpret_cast<VoxelBuffer3D*>(world_); }
    Bounds3Df* GetVoxelBufferDataBounds() { return reinterpret_cast<Bounds3Df*>(world_); }
    int GetFactor() const { return factor; }
    void SetFactor(int f) { factor = f; }
    vx_tracer* Handle() { return tracer_; }
    void Free() {
        if (tracer_) { vx_tracer_destroy(tracer_); tracer_ = nullptr; }
        resultsCPU = RayTraceResults<float3>(0);
    }
    RayTraceResults<float3> Raytrace(std::vector<float3> origin, std::vector<float3> ray) { // VolumeRaytracer.cu:574-618
        auto& r = resultsCPU;
        const size_t n = origin.size();
        vx_hits h{};
        h.hit = hit_.data(); h.pos = &r.hitPoint[0].x; h.normal = &r.normal[0].x; h.steps = r.steps.get();
        h.voxel = r.voxel.get(); h.brick = r.brick.get();
        vxrt_check(vx_trace(tracer_, &origin[0].x, &ray[0].x, n, nullptr, &h), "vx_trace");
        vx_hits_postprocess(&origin[0].x, &r.hitPoint[0].x, n, reinterpret_cast<uint8_t*>(r.valid.get()), r.distance.get());
        for (size_t i = 0; i < n; ++i) // the reference's voxelIndex is a float expression on coarse dims (unusable); here: brick index
            r.voxelIndex[i] = hit_[i] ? (int)r.brick[i] : -1;
        return r;
    }
};

// VolumeRaytracer.cuh:379-516.  Builds the brickmap on the device; factor must be a multiple of 8.
static std::tuple<VoxelBuffer3D, VoxelBuffer3D*, Bounds3Df*> GenerateLowresVoxelBuffer(const VoxelBuffer3D& originalData, int factor = 8) {
    int dev = 0;
    cudaGetDevice(&dev);
    const uint32_t X = originalData.dimensions[0] ? originalData.dimensions[0] : originalData.full_dims[0];
    const uint32_t Y = originalData.dimensions[1] ? originalData.dimensions[1] : originalData.full_dims[1];
    const uint32_t Z = originalData.dimensions[2] ? originalData.dimensions[2] : originalData.full_dims[2];
    vx_world* w = nullptr;
    if (originalData.from_noise) vxrt_check(vx_world_build_noise(&w, X, Y, Z, factor, 0, 0, 0, dev), "vx_world_build_noise");
    else vxrt_check(vx_world_from_dense_bits(&w, originalData.grid.data, X, Y, Z, factor, dev), "vx_world_from_dense_bits");
    VoxelBuffer3D low;
    low.world = w;
    low.dimensions[0] = (uint16_t)(X / factor); low.dimensions[1] = (uint16_t)(Y / factor); low.dimensions[2] = (uint16_t)(Z / factor);
    return std::make_tuple(low, new VoxelBuffer3D[1], new Bounds3Df[1]);
}

namespace Graphics { // Renderer.cuh:25-56
struct BGRA8888 { uint8_t b, g, r, a; };
struct Environment { float3 LightDirection, LightColor, AmbientColor; };
struct ShimState {
    vx_env env{{0.57735026f, 0.57735026f, 0.57735026f}, {2, 2, 2}, {0.5f, 0.5f, 0.5f}};
    float fov = 90.0f;
    float2 ortho{10, 10};
    uint32_t frame = 0;
    // as shipped: checkerboard on, DEBUG_VIEW on, shadow call commented out (Renderer.cu:4-5,102);
    // define VXRT_COMPAT_SHADED to get calculateColor + Tonemap with the shadow ray restored
#ifdef VXRT_COMPAT_SHADED
    int debug_view = 0, shadows = 1;
#else
    int debug_view = 1, shadows = 0;
#endif
    int checkerboard = 1;
};
inline ShimState& State() { static ShimState s; return s; }
inline void GetDirections(float3 eularAngles, float3* forwad, float3* up, float3* right) {
    vx_get_directions(&eularAngles.x, &forwad->x, &up->x, &right->x);
}
inline void SetEnvironment(const Environment& e) {
    vx_env& v = State().env;
    memcpy(v.light_dir, &e.LightDirection, 12); memcpy(v.light_color, &e.LightColor, 12); memcpy(v.ambient, &e.AmbientColor, 12);
}
inline void SetFOV(float fov) { State().fov = fov; }
inline void SetOrthoWindowSize(float2 s) { State().ortho = s; }
inline void RenderScreen(VoxelRaytracer3D* rt, uint32_t screen_width, uint32_t screen_height, void* d_screen_texture, float3 origin,
                         float3 camera_fwd, float3 camera_up, float3 camera_right) { // Renderer.cu:305-328
    ShimState& s = State();
    vx_camera cam{};
    memcpy(cam.origin, &origin, 12); memcpy(cam.fwd, &camera_fwd, 12); memcpy(cam.up, &camera_up, 12); memcpy(cam.right, &camera_right, 12);
    cam.fov_deg = s.fov;
#ifdef ORTHO // the reference's compile-time switch (Renderer.cuh:13, Renderer.cu:202-204)
    cam.ortho_w = s.ortho.x; cam.ortho_h = s.ortho.y;
#endif
    vx_render_opts o{};
    o.checkerboard = s.checkerboard; o.frame_number = s.frame; o.debug_view = s.debug_view; o.shadows = s.shadows; o.crosshair = 1;
    o.norm_mode = VX_NORM_DEVICE;
    s.frame++; // hFrameInfo.FrameNumber++ after the copy to the device (Renderer.cu:310,322)
    vxrt_check(vx_render(rt->Handle(), &cam, &s.env, &o, d_screen_texture, nullptr, screen_width, screen_height, nullptr), "vx_render");
    cudaError_t err = cudaDeviceSynchronize();
    if (err != cudaSuccess) { fprintf(stderr, "CUDA error: %s\n", cudaGetErrorString(err)); exit(EXIT_FAILURE); }
}
} // namespace Graphics
} // namespace GPUDDA

// VoxelWorldBuilder.cuh:12-32.  The noise world is generated on the device inside GenerateLowresVoxelBuffer;
// the returned buffer only records the size (grid.Raw() is a one-word allocation so `delete[]` stays valid).
static GPUDDA::VoxelBuffer3D CreateVoxels(uint3 size) {
    GPUDDA::VoxelBuffer3D v;
    v.dimensions[0] = (uint16_t)size.x; v.dimensions[1] = (uint16_t)size.y; v.dimensions[2] = (uint16_t)size.z;
    v.full_dims[0] = size.x; v.full_dims[1] = size.y; v.full_dims[2] = size.z;
    v.grid = GPUDDA::BitArray(32, false);
    v.from_noise = true;
    return v;
}
