// headless_main.cu — the reference's VoxelApp/main.cu:18-66,165-167 with the SDL window removed, compiled
// against include/voxelrt_compat/ and linked with libvxrt.so.  Everything between the markers is the
// reference's call sequence with its own names; only the presentation loop differs (frames are rendered into
// the cudaMalloc'd buffer and copied back, as main.cu:165-167 does into the SDL texture).
//   nvcc -std=c++17 -Iinclude/voxelrt_compat examples/headless_main.cu -Lvoxelengine_b200/lib -lvxrt -o headless
#include "VolumeRaytracer.cuh"
#include "Renderer.cuh"
#include "VoxelWorldBuilder.cuh"
#include <chrono>
#include <fstream>
#include <iostream>

using namespace GPUDDA::Graphics;
using namespace GPUDDA;

constexpr uint32_t width = 1280;
constexpr uint32_t height = 720;

int main(int argc, char** argv) {
    int factor = argc > 1 ? atoi(argv[1]) : 32;
    int frames = argc > 2 ? atoi(argv[2]) : 4;
    auto t0 = std::chrono::high_resolution_clock::now();
    auto buffer = CreateVoxels(make_uint3(1024, 1024, 1024));
    auto buffers = GenerateLowresVoxelBuffer(buffer, factor);
    auto t3 = std::chrono::high_resolution_clock::now();
    std::cout << "Voxel + buffer generation time: " << std::chrono::duration_cast<std::chrono::milliseconds>(t3 - t0).count() << "ms" << std::endl;
    delete[] buffer.grid.Raw();

    VoxelRaytracer3D* raytracer = new GPUDDA::VoxelRaytracer3D(1);
    auto low_res_buffer = std::get<0>(buffers);
    auto low_res_grid_data = std::get<1>(buffers);
    auto bounds = std::get<2>(buffers);
    auto count = low_res_buffer.dimensions[0] * low_res_buffer.dimensions[1] * low_res_buffer.dimensions[2];
    raytracer->UploadVoxelBuffer(low_res_buffer);
    raytracer->UploadVoxelBufferDatas(low_res_grid_data, count);
    raytracer->UploadVoxelBufferDataBounds(bounds, count);
    raytracer->SetFactor(factor);

    void* d_pixels;
    float3 cam_pos = {256, 256, 256};
    float3 cam_up = {0, 1, 0}, cam_right = {1, 0, 0}, cam_forward = {0, 0, 1}, cam_eular = {0, 0, 0};
    Graphics::Environment env;
    env.LightDirection = {0.57735026f, 0.57735026f, 0.57735026f}; // normalize({1,1,1})
    env.LightColor = {2, 2, 2};
    env.AmbientColor = {0.5f, 0.5f, 0.5f};
    SetEnvironment(env);
    SetFOV(90);
    SetOrthoWindowSize(make_float2(10, 10));
    cudaMalloc(&d_pixels, width * height * 4);
    cudaMemset(d_pixels, 255, width * height * 4);
    std::vector<uint8_t> pixels(width * height * 4);
    for (int f = 0; f < frames; ++f) {
        cam_eular.y += 0.05f; // stands in for the mouse-look input of main.cu:72-161
        auto f0 = std::chrono::high_resolution_clock::now();
        GetDirections(cam_eular, &cam_forward, &cam_up, &cam_right);
        Graphics::RenderScreen(raytracer, width, height, d_pixels, cam_pos, cam_forward, cam_up, cam_right);
        cudaMemcpy(pixels.data(), d_pixels, pixels.size(), cudaMemcpyDeviceToHost);
        auto f1 = std::chrono::high_resolution_clock::now();
        std::cout << "frame " << f << ": " << std::chrono::duration<double, std::milli>(f1 - f0).count() << " ms" << std::endl;
    }
    std::ofstream("frame.bgra", std::ios::binary).write((const char*)pixels.data(), pixels.size());
    delete raytracer;
    return 0;
}
