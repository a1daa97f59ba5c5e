// multi_gpu_main.cu — the reference's frame loop (VoxelApp/main.cu:70-190: RenderScreen into a cudaMalloc'd buffer, cudaMemcpy
// to the host every frame) on N GPUs of one box, in ONE process and without Python: plain C++ over the C ABI (include/vxrt.h).
//
//   * the world is replicated: every device builds it itself (deterministic), one host thread per device;
//   * the frame lives on the first device; every other device gets peer access to it (NVLink / NVSwitch) and all devices
//     render their share of the frame's 8-row bands (vx_render_opts::band_count / band_index) straight into it — the
//     exchange step of the partition is the kernels' own pixel stores, there is no gather;
//   * one event per device closes the frame, then a single cudaMemcpy brings it down, as main.cu:165-167 does.
// The first frame is also rendered whole on the first device alone and compared byte for byte.
//
//   nvcc -std=c++17 -arch=sm_100a -Iinclude examples/multi_gpu_main.cu -Lvoxelengine_b200/lib -lvxrt -o examples/multi_gpu
//   examples/multi_gpu [--devices 0,1,2,3] [--world X Y Z] [--frame W H] [--frames K] [--out frame.bgra]
// (--devices may name a device more than once: "0,0" runs two band sets on one GPU, which is how the single-GPU test drives it.)
#include "vxrt.h"
#include <cuda_runtime.h>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <thread>
#include <vector>

#define CHECK_VX(call) do { int rc_ = (call); if (rc_ != VX_OK) { fprintf(stderr, "%s failed (%d): %s\n", #call, rc_, vx_last_error()); exit(1); } } while (0)
#define CHECK_CUDA(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { fprintf(stderr, "%s failed: %s\n", #call, cudaGetErrorString(e_)); exit(1); } } while (0)

static uint32_t crc32_of(const uint8_t* p, size_t n) {
    static uint32_t table[256];
    static bool init = false;
    if (!init) {
        for (uint32_t i = 0; i < 256; ++i) { uint32_t c = i; for (int k = 0; k < 8; ++k) c = (c & 1u) ? 0xEDB88320u ^ (c >> 1) : c >> 1; table[i] = c; }
        init = true;
    }
    uint32_t c = 0xFFFFFFFFu;
    for (size_t i = 0; i < n; ++i) c = table[(c ^ p[i]) & 255u] ^ (c >> 8);
    return c ^ 0xFFFFFFFFu;
}

struct Device {
    int id = 0;
    vx_world* world = nullptr;
    vx_tracer* tracer = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;
};

int main(int argc, char** argv) {
    std::vector<int> ids;
    uint32_t X = 1024, Y = 256, Z = 1024, W = 1920, H = 1080;
    int frames = 8;
    std::string out_path;
    for (int i = 1; i < argc; ++i) {
        if (!strcmp(argv[i], "--devices") && i + 1 < argc) { for (char* t = strtok(argv[++i], ","); t; t = strtok(nullptr, ",")) ids.push_back(atoi(t)); }
        else if (!strcmp(argv[i], "--world") && i + 3 < argc) { X = atoi(argv[i + 1]); Y = atoi(argv[i + 2]); Z = atoi(argv[i + 3]); i += 3; }
        else if (!strcmp(argv[i], "--frame") && i + 2 < argc) { W = atoi(argv[i + 1]); H = atoi(argv[i + 2]); i += 2; }
        else if (!strcmp(argv[i], "--frames") && i + 1 < argc) frames = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--out") && i + 1 < argc) out_path = argv[++i];
        else { fprintf(stderr, "unknown argument %s\n", argv[i]); return 2; }
    }
    if (ids.empty()) { int n = 0; CHECK_VX(vx_device_count(&n)); for (int d = 0; d < n; ++d) ids.push_back(d); }
    if (ids.empty()) { fprintf(stderr, "no CUDA device: there is no CPU fallback\n"); return 1; }
    const uint32_t N = (uint32_t)ids.size();
    std::vector<Device> dev(N);

    // ---- start-up (main.cu:18-66): every device builds its replica of the world, in parallel ----
    auto t0 = std::chrono::high_resolution_clock::now();
    {
        std::vector<std::thread> pool;
        for (uint32_t k = 0; k < N; ++k)
            pool.emplace_back([&, k] {
                dev[k].id = ids[k];
                CHECK_VX(vx_world_build_noise(&dev[k].world, X, Y, Z, 8, 0, 0, 0, ids[k]));
                CHECK_VX(vx_tracer_create(&dev[k].tracer, dev[k].world, 16));
                CHECK_CUDA(cudaSetDevice(ids[k]));
                CHECK_CUDA(cudaStreamCreateWithFlags(&dev[k].stream, cudaStreamNonBlocking));
                CHECK_CUDA(cudaEventCreateWithFlags(&dev[k].done, cudaEventDisableTiming));
                if (ids[k] != ids[0]) { // the frame is in the first device's memory
                    int can = 0;
                    CHECK_CUDA(cudaDeviceCanAccessPeer(&can, ids[k], ids[0]));
                    if (!can) { fprintf(stderr, "device %d cannot access device %d's memory\n", ids[k], ids[0]); exit(1); }
                    cudaError_t e = cudaDeviceEnablePeerAccess(ids[0], 0);
                    if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) CHECK_CUDA(e);
                    cudaGetLastError();
                    // a frame in another GPU's memory is shaded inside the sun-ray march: its pixel stores spread over the kernel
                    if (N >= 4) CHECK_VX(vx_tracer_set_fused_shade(dev[k].tracer, 1));
                }
            });
        for (auto& t : pool) t.join();
    }
    vx_world_info info;
    CHECK_VX(vx_world_get_info(dev[0].world, &info));
    printf("devices=%u world=%ux%ux%u occupied_bricks=%llu build_ms=%.1f startup_ms=%.1f\n", N, X, Y, Z, (unsigned long long)info.occupied_bricks,
           info.build_ms, std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - t0).count());

    // camera: 16 voxels above the terrain at the middle of the +z edge, looking across the map and down (SURVEY.md 8d rule)
    vx_camera cam{};
    {
        const float o[3] = {X * 0.5f + 0.5f, Y + 10.0f, Z - 64.0f + 0.5f}, d[3] = {0.0f, -1.0f, 0.0f};
        uint8_t hit = 0; int32_t voxel[3] = {0, 0, 0};
        vx_hits h{}; h.hit = &hit; h.voxel = voxel;
        CHECK_VX(vx_trace(dev[0].tracer, o, d, 1, nullptr, &h));
        const int top = hit ? voxel[1] : -1;
        cam.origin[0] = X * 0.5f; cam.origin[1] = (float)(top < (int)Y - 1 ? top + 16 : Y + 32); cam.origin[2] = Z - 64.0f;
    }
    float euler[3] = {-0.45f, 0.0f, 0.0f};
    cam.fov_deg = 90.0f;
    const float s3 = 0.57735026f;
    vx_env env = {{s3, s3, s3}, {2.0f, 2.0f, 2.0f}, {0.5f, 0.5f, 0.5f}}; // main.cu:58-62

    CHECK_CUDA(cudaSetDevice(ids[0]));
    const size_t bytes = (size_t)W * H * 4;
    void *d_frame = nullptr, *d_whole = nullptr;
    CHECK_CUDA(cudaMalloc(&d_frame, bytes));
    CHECK_CUDA(cudaMalloc(&d_whole, bytes));
    CHECK_CUDA(cudaMemset(d_frame, 255, bytes)); // main.cu:69
    CHECK_CUDA(cudaMemset(d_whole, 255, bytes));
    uint8_t* pixels = nullptr;
    CHECK_CUDA(cudaMallocHost(&pixels, bytes));
    std::vector<uint8_t> whole(bytes);

    bool identical = true;
    double sum_ms = 0.0;
    int timed = 0;
    for (int f = 0; f < frames; ++f) {
        euler[1] += 0.01f; // stands in for the mouse look of main.cu:72-161
        vx_get_directions(euler, cam.fwd, cam.up, cam.right);
        vx_render_opts o{};
        o.checkerboard = 0; o.frame_number = (uint32_t)f; o.debug_view = 0; o.shadows = 1; o.crosshair = 1; o.norm_mode = VX_NORM_DEVICE;
        auto f0 = std::chrono::high_resolution_clock::now();
        for (uint32_t k = 0; k < N; ++k) { // every device renders its bands into the one frame
            vx_render_opts ok = o;
            ok.band_count = N > 1 ? N : 0; ok.band_index = N > 1 ? k : 0;
            CHECK_VX(vx_render(dev[k].tracer, &cam, &env, &ok, d_frame, nullptr, W, H, dev[k].stream));
            CHECK_CUDA(cudaSetDevice(dev[k].id));
            CHECK_CUDA(cudaEventRecord(dev[k].done, dev[k].stream));
        }
        CHECK_CUDA(cudaSetDevice(ids[0]));
        for (uint32_t k = 0; k < N; ++k) CHECK_CUDA(cudaStreamWaitEvent(dev[0].stream, dev[k].done, 0)); // a kernel's peer stores are done when it is
        CHECK_CUDA(cudaMemcpyAsync(pixels, d_frame, bytes, cudaMemcpyDeviceToHost, dev[0].stream));
        CHECK_CUDA(cudaStreamSynchronize(dev[0].stream));
        const double ms = std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now() - f0).count();
        if (f >= 2) { sum_ms += ms; ++timed; }
        if (f == 0) { // the same frame rendered whole on the first device
            CHECK_VX(vx_tracer_set_fused_shade(dev[0].tracer, 0));
            CHECK_VX(vx_render(dev[0].tracer, &cam, &env, &o, d_whole, nullptr, W, H, dev[0].stream));
            CHECK_CUDA(cudaMemcpyAsync(whole.data(), d_whole, bytes, cudaMemcpyDeviceToHost, dev[0].stream));
            CHECK_CUDA(cudaStreamSynchronize(dev[0].stream));
            identical = memcmp(whole.data(), pixels, bytes) == 0;
            printf("frame0_crc32=%08x whole_frame_crc32=%08x frame_identical=%d\n", crc32_of(pixels, bytes), crc32_of(whole.data(), bytes), identical ? 1 : 0);
        }
        printf("frame %d: %.3f ms\n", f, ms);
    }
    if (timed) printf("mean_frame_ms=%.3f (frames 2..%d, render on %u devices + copy to the host)\n", sum_ms / timed, frames - 1, N);
    if (!out_path.empty()) { FILE* g = fopen(out_path.c_str(), "wb"); if (g) { fwrite(pixels, 1, bytes, g); fclose(g); } }
    for (uint32_t k = 0; k < N; ++k) {
        CHECK_CUDA(cudaSetDevice(dev[k].id));
        cudaStreamSynchronize(dev[k].stream);
        vx_tracer_destroy(dev[k].tracer); vx_world_destroy(dev[k].world);
        cudaEventDestroy(dev[k].done); cudaStreamDestroy(dev[k].stream);
    }
    cudaSetDevice(ids[0]);
    cudaFree(d_frame); cudaFree(d_whole); cudaFreeHost(pixels);
    return identical ? 0 : 3;
}
