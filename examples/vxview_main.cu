// vxview_main.cu — the reference's interactive loop (VoxelApp/main.cu:18-195) run headless: the SDL window, event pump and
// presenter are replaced by an input script (examples/input_replay.h, one line per frame with the keys held, the mouse
// position / button and the wheel), frames are rendered through the reference's own API names (include/voxelrt_compat)
// into the cudaMalloc'd BGRA buffer and copied back as main.cu:165-167 does for the SDL texture, selected frames are written
// as PPM, and the frame time is logged with the reference's running average (main.cu:173-183).
//
// The reference renders every other row of every other column per frame (Renderer.cu:186-194) into a buffer that persists,
// so an image fills in over two frames: the dumps show exactly that.
//
//   nvcc -std=c++17 -Iinclude/voxelrt_compat examples/vxview_main.cu -Lvoxelengine_b200/lib -lvxrt -o vxview
//   vxview [--size X Y Z] [--factor F] [--frames N] [--script file] [--dump every] [--out prefix] [--width W --height H]
//   Without a script: hold W with shift, yaw by 10 px per frame with the button down (a fly-over).
#include "VolumeRaytracer.cuh"
#include "Renderer.cuh"
#include "VoxelWorldBuilder.cuh"
#include "input_replay.h"
#include <chrono>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

using namespace GPUDDA::Graphics;
using namespace GPUDDA;

int main(int argc, char** argv) {
    uint32_t width = 1280, height = 720, X = 1024, Y = 1024, Z = 1024; // main.cu:15-16, :34
    int factor = 32, frames = 8, dump = 0;
    std::string script, out = "frame";
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&](int k) { return i + k < argc ? atoi(argv[i + k]) : 0; };
        if (a == "--size" && i + 3 < argc) { X = next(1); Y = next(2); Z = next(3); i += 3; }
        else if (a == "--factor") { factor = next(1); ++i; }
        else if (a == "--frames") { frames = next(1); ++i; }
        else if (a == "--dump") { dump = next(1); ++i; }
        else if (a == "--width") { width = next(1); ++i; }
        else if (a == "--height") { height = next(1); ++i; }
        else if (a == "--script" && i + 1 < argc) { script = argv[++i]; }
        else if (a == "--out" && i + 1 < argc) { out = argv[++i]; }
        else { std::cerr << "unknown argument " << a << std::endl; return 2; }
    }
    std::vector<replay::InputFrame> inputs;
    if (!script.empty()) {
        std::ifstream in(script);
        if (!in) { std::cerr << "cannot read " << script << std::endl; return 2; }
        std::string line;
        while (std::getline(in, line)) {
            if (line.empty() || line[0] == '#') continue;
            replay::InputFrame f;
            if (!replay::parse_line(line.c_str(), f)) { std::cerr << "bad script line: " << line << std::endl; return 2; }
            inputs.push_back(f);
        }
        frames = (int)inputs.size();
    } else {
        for (int f = 0; f < frames; ++f) {
            replay::InputFrame in;
            in.w = true; in.shift = true; in.click = f > 0; in.mouse_x = 640 + 10 * f; in.mouse_y = 360;
            inputs.push_back(in);
        }
    }

    // ---- main.cu:18-66 with its own names ----
    auto t0 = std::chrono::high_resolution_clock::now();
    auto buffer = CreateVoxels(make_uint3(X, Y, Z));
    auto buffers = GenerateLowresVoxelBuffer(buffer, factor);
    auto t3 = std::chrono::high_resolution_clock::now();
    std::cout << "Voxel + buffer generation time: " << std::chrono::duration_cast<std::chrono::milliseconds>(t3 - t0).count() << "ms" << std::endl;
    delete[] buffer.grid.Raw();
    VoxelRaytracer3D* raytracer = new GPUDDA::VoxelRaytracer3D(1);
    auto low_res_buffer = std::get<0>(buffers);
    auto count = low_res_buffer.dimensions[0] * low_res_buffer.dimensions[1] * low_res_buffer.dimensions[2];
    raytracer->UploadVoxelBuffer(low_res_buffer);
    raytracer->UploadVoxelBufferDatas(std::get<1>(buffers), count);
    raytracer->UploadVoxelBufferDataBounds(std::get<2>(buffers), count);
    raytracer->SetFactor(factor);

    void* d_pixels;
    replay::CameraRig rig; // cam_pos {256,256,256}, eular 0, ortho window 10 x 10
    Graphics::Environment env;
    env.LightDirection = {0.57735026f, 0.57735026f, 0.57735026f}; // normalize({1,1,1})
    env.LightColor = {2, 2, 2};
    env.AmbientColor = {0.5f, 0.5f, 0.5f};
    SetEnvironment(env);
    SetFOV(90);
    SetOrthoWindowSize(rig.ortho);
    cudaMalloc(&d_pixels, (size_t)width * height * 4);
    cudaMemset(d_pixels, 255, (size_t)width * height * 4);
    std::vector<uint8_t> pixels((size_t)width * height * 4);

    // ---- the render callback (main.cu:71-168) per scripted frame, and the timing of the outer loop (:170-190) ----
    replay::FrameClock clock;
    for (int f = 0; f < frames; ++f) {
        auto f0 = std::chrono::high_resolution_clock::now();
        replay::apply_input(rig, inputs[f]);
        SetOrthoWindowSize(rig.ortho);
        GetDirections(rig.eular, &rig.forward, &rig.up, &rig.right);
        RenderScreen(raytracer, width, height, d_pixels, rig.pos, rig.forward, rig.up, rig.right);
        cudaMemcpy(pixels.data(), d_pixels, pixels.size(), cudaMemcpyDeviceToHost);
        auto f1 = std::chrono::high_resolution_clock::now();
        const double td = std::chrono::duration_cast<std::chrono::microseconds>(f1 - f0).count() / 1000.0;
        clock.add(td);
        printf("frame %d: %.3f ms, pos %.3f %.3f %.3f, eular %.4f %.4f, Avg FPS: %.2f\n", f, td, rig.pos.x, rig.pos.y, rig.pos.z,
               rig.eular.x, rig.eular.y, clock.avg_fps());
        if (dump > 0 && (f % dump == dump - 1 || f == frames - 1)) {
            char name[512];
            snprintf(name, sizeof name, "%s_%04d.ppm", out.c_str(), f);
            if (!replay::write_ppm(name, pixels.data(), width, height)) { std::cerr << "cannot write " << name << std::endl; return 1; }
        }
    }
    cudaFree(d_pixels);
    delete raytracer;
    return 0;
}
