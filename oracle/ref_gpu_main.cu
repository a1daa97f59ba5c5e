/*
 * ref_gpu_main.cu — headless driver around the UNMODIFIED reference (test infrastructure).
 * Linked by oracle/Makefile with /root/reference/VoxelRT/{VolumeRaytracer,Renderer,VoxelWorldBuilder,
 * cuda_noise}.cu into oracle/_ref/ref_gpu (default flags) and ref_gpu_nofma (-fmad=false, the parity
 * authority of SURVEY.md §8c).  It performs the start-up sequence of VoxelApp/main.cu:18-66 without
 * SDL and then one of:
 *   build  X Y Z factor out.bin                      dump GenerateLowresVoxelBuffer(CreateVoxels(..))
 *   trace  X Y Z factor rays.bin out.bin             VoxelRaytracer3D::Raytrace on rays.bin (n, origins, dirs)
 *   render X Y Z factor W H px py pz ex ey ez frames out.bgra [ow oh]   Graphics::RenderScreen, `frames` times
 *                                                              (ow oh: SetOrthoWindowSize, used by the -DORTHO build);
 *                                                              out.bgra = the buffer after frame 0; with VXREF_DUMP_LAST=path
 *                                                              the buffer after the last frame goes to `path` (the checkerboard
 *                                                              halves of consecutive frames fill it, main.cu:165-167)
 * Timings are printed as "key=value" lines.
 */
#include "VolumeRaytracer.cuh"
#include "Renderer.cuh"
#include "VoxelWorldBuilder.cuh"
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

using namespace GPUDDA;
using namespace GPUDDA::Graphics;

static double now_ms() {
    return std::chrono::duration<double, std::milli>(std::chrono::high_resolution_clock::now().time_since_epoch()).count();
}

struct RefWorld {
    VoxelBuffer3D dense, top;
    VoxelBuffer3D* bricks;
    Bounds3Df* bounds;
    size_t count;
    int factor;
};

static RefWorld make_world(int X, int Y, int Z, int factor) {
    RefWorld w;
    double t0 = now_ms();
    w.dense = CreateVoxels(make_uint3(X, Y, Z));
    double t1 = now_ms();
    auto t = GenerateLowresVoxelBuffer(w.dense, factor);
    double t2 = now_ms();
    w.top = std::get<0>(t); w.bricks = std::get<1>(t); w.bounds = std::get<2>(t);
    w.count = (size_t)w.top.dimensions[0] * w.top.dimensions[1] * w.top.dimensions[2];
    w.factor = factor;
    printf("ref_voxel_generation_ms=%.3f\nref_buffer_generation_ms=%.3f\n", t1 - t0, t2 - t1);
    return w;
}

static VoxelRaytracer3D* upload(RefWorld& w, size_t rays) {
    double t0 = now_ms();
    VoxelRaytracer3D* rt = new VoxelRaytracer3D(rays);
    rt->UploadVoxelBuffer(w.top);
    rt->UploadVoxelBufferDatas(w.bricks, w.count);
    rt->UploadVoxelBufferDataBounds(w.bounds, w.count);
    rt->SetFactor(w.factor);
    cudaDeviceSynchronize();
    printf("ref_upload_ms=%.3f\n", now_ms() - t0);
    return rt;
}

int main(int argc, char** argv) {
    if (argc < 6) { fprintf(stderr, "usage: see header comment\n"); return 2; }
    const char* cmd = argv[1];
    const int X = atoi(argv[2]), Y = atoi(argv[3]), Z = atoi(argv[4]), factor = atoi(argv[5]);
    RefWorld w = make_world(X, Y, Z, factor);
    if (!strcmp(cmd, "build")) {
        FILE* f = fopen(argv[6], "wb");
        const size_t wpb = (size_t)factor * factor * factor / 32;
        uint64_t hdr[2] = {w.count, wpb};
        fwrite(hdr, 8, 2, f);
        fwrite(w.top.grid.Raw(), 4, (w.count + 31) / 32, f);
        std::vector<uint32_t> zero(wpb, 0u);
        for (size_t i = 0; i < w.count; ++i)
            fwrite(w.bricks[i].dimensions[0] ? w.bricks[i].grid.Raw() : zero.data(), 4, wpb, f);
        fwrite(w.bounds, sizeof(Bounds3Df), w.count, f);
        fclose(f);
        return 0;
    }
    if (!strcmp(cmd, "trace")) {
        FILE* f = fopen(argv[6], "rb");
        uint64_t n = 0;
        if (!f || fread(&n, 8, 1, f) != 1) { fprintf(stderr, "bad rays file\n"); return 2; }
        std::vector<float3> o(n), d(n);
        if (fread(o.data(), 12, n, f) != n || fread(d.data(), 12, n, f) != n) { fprintf(stderr, "short rays file\n"); return 2; }
        fclose(f);
        VoxelRaytracer3D* rt = upload(w, n);
        auto r = rt->Raytrace(o, d); // prints "Raytracing time: .. ms"
        r = rt->Raytrace(o, d);      // second call: warm
        FILE* g = fopen(argv[7], "wb");
        fwrite(&n, 8, 1, g);
        fwrite(r.hitPoint.get(), 12, n, g);
        fwrite(r.normal.get(), 12, n, g);
        fwrite(r.steps.get(), 4, n, g);
        fclose(g);
        return 0;
    }
    if (!strcmp(cmd, "render")) {
        if (argc < 16) { fprintf(stderr, "render: missing arguments\n"); return 2; }
        const uint32_t W = atoi(argv[6]), H = atoi(argv[7]);
        float3 pos = make_float3(atof(argv[8]), atof(argv[9]), atof(argv[10]));
        float3 eul = make_float3(atof(argv[11]), atof(argv[12]), atof(argv[13]));
        const int frames = atoi(argv[14]);
        VoxelRaytracer3D* rt = upload(w, 1);
        Environment env; // VoxelApp/main.cu:58-66
        env.LightDirection = normalize(make_float3(1, 1, 1));
        env.LightColor = make_float3(2, 2, 2);
        env.AmbientColor = make_float3(0.5f, 0.5f, 0.5f);
        SetEnvironment(env);
        SetFOV(90);
        SetOrthoWindowSize(argc >= 18 ? make_float2(atof(argv[16]), atof(argv[17])) : make_float2(10, 10));
        void* d_pixels = nullptr;
        cudaMalloc(&d_pixels, (size_t)W * H * 4);
        cudaMemset(d_pixels, 255, (size_t)W * H * 4); // main.cu:69
        float3 fwd, up, right;
        GetDirections(eul, &fwd, &up, &right);
        printf("ref_basis=%.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g %.9g\n", fwd.x, fwd.y, fwd.z, up.x, up.y, up.z, right.x, right.y, right.z);
        cudaEvent_t e0, e1;
        cudaEventCreate(&e0); cudaEventCreate(&e1);
        std::vector<uint8_t> first((size_t)W * H * 4);
        for (int i = 0; i < frames; ++i) {
            cudaEventRecord(e0);
            RenderScreen(rt, W, H, d_pixels, pos, fwd, up, right);
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms = 0;
            cudaEventElapsedTime(&ms, e0, e1);
            printf("ref_frame_ms[%d]=%.4f\n", i, ms);
            if (i == 0) cudaMemcpy(first.data(), d_pixels, first.size(), cudaMemcpyDeviceToHost);
        }
        FILE* g = fopen(argv[15], "wb");
        fwrite(first.data(), 1, first.size(), g); // frame 0 alone (FrameNumber 0 during the launch)
        fclose(g);
        if (const char* last = getenv("VXREF_DUMP_LAST")) {
            cudaMemcpy(first.data(), d_pixels, first.size(), cudaMemcpyDeviceToHost);
            FILE* l = fopen(last, "wb");
            fwrite(first.data(), 1, first.size(), l);
            fclose(l);
        }
        return 0;
    }
    fprintf(stderr, "unknown command %s\n", cmd);
    return 2;
}
