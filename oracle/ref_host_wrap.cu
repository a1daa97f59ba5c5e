/*
 * ref_host_wrap.cu — C entry points around the REFERENCE's own code compiled for the host
 * (test infrastructure; built only into oracle/_ref/, never shipped, never linked by the product).
 * Calls GPUDDA::GenerateLowresVoxelBuffer (VolumeRaytracer.cuh:379), GPUDDA::Raytrace
 * (VolumeRaytracer.cu:354), DDARayTraversal (:176) and RayIntersectsAABB (:124) as they are.
 */
#include "VolumeRaytracer.cuh"
#include "helper_math.h"
#include <cstring>
#include <thread>
#include <vector>
using namespace GPUDDA;

struct vxref_world {
    VoxelBuffer3D dense;
    VoxelBuffer3D top;
    VoxelBuffer3D* bricks;
    Bounds3Df* bounds;
    int factor;
    size_t nbricks;
};

extern "C" vxref_world* vxref_build(const uint32_t* bits, uint32_t X, uint32_t Y, uint32_t Z, int factor) {
    vxref_world* w = new vxref_world();
    size_t nvox = (size_t)X * Y * Z;
    w->dense.grid = BitArray(nvox, false);
    std::memcpy(w->dense.grid.Raw(), bits, (nvox + 31) / 32 * 4);
    w->dense.dimensions[0] = X; w->dense.dimensions[1] = Y; w->dense.dimensions[2] = Z;
    auto t = GenerateLowresVoxelBuffer(w->dense, factor);
    w->top = std::get<0>(t); w->bricks = std::get<1>(t); w->bounds = std::get<2>(t);
    w->factor = factor;
    w->nbricks = (size_t)w->top.dimensions[0] * w->top.dimensions[1] * w->top.dimensions[2];
    return w;
}
/* A reference-form world assembled from brickmap arrays the caller keeps alive (top-level bits in the tiled index, factor^3/32
 * words per brick in tiled coarse order, six floats of bounds per brick): what GenerateLowresVoxelBuffer returns, without the
 * dense grid it needs as input — 2^35 voxels at 8192x512x8192, beyond the reference's 32-bit voxel index.  Raytrace itself only
 * indexes the coarse grid and single bricks, so the reference's own march runs on such a world unchanged.  BitArray keeps its
 * two fields private (VolumeRaytracer.cuh:204-223: size_t size; uint32_t* data), hence the layout-checked memcpy. */
struct RawBits { size_t size; uint32_t* data; };
static_assert(sizeof(RawBits) == sizeof(BitArray), "BitArray layout");
static void adopt(BitArray& dst, size_t bits, const uint32_t* words) {
    RawBits r{bits, const_cast<uint32_t*>(words)};
    std::memcpy((void*)&dst, &r, sizeof r);
}
extern "C" vxref_world* vxref_from_brickmap(const uint32_t* top_words, const uint32_t* brick_words, const float* bounds6,
                                            uint32_t X, uint32_t Y, uint32_t Z, int factor) {
    static_assert(sizeof(Bounds3Df) == 24, "Bounds3Df layout");
    vxref_world* w = new vxref_world();
    const uint32_t cx = X / factor, cy = Y / factor, cz = Z / factor;
    w->factor = factor;
    w->nbricks = (size_t)cx * cy * cz;
    adopt(w->top.grid, w->nbricks, top_words);
    w->top.dimensions[0] = (uint16_t)cx; w->top.dimensions[1] = (uint16_t)cy; w->top.dimensions[2] = (uint16_t)cz;
    const size_t bits = (size_t)factor * factor * factor, wpb = bits / 32;
    w->bricks = new VoxelBuffer3D[w->nbricks];
    w->bounds = reinterpret_cast<Bounds3Df*>(const_cast<float*>(bounds6));
    for (size_t i = 0; i < w->nbricks; ++i) {
        const bool occupied = (top_words[i >> 5] >> (i & 31)) & 1u;
        if (!occupied) continue; /* empty brick: dimensions 0, no grid (VolumeRaytracer.cuh:379-516) */
        adopt(w->bricks[i].grid, bits, brick_words + i * wpb);
        w->bricks[i].dimensions[0] = w->bricks[i].dimensions[1] = w->bricks[i].dimensions[2] = (uint16_t)factor;
    }
    return w;
}
extern "C" void vxref_free_view(vxref_world* w) { if (w) { delete[] w->bricks; delete w; } }
/* helper_math.h normalize (:1325-1329, host rsqrtf here) and reflect (:1427-1430) as the reference's AO sample loop uses them */
extern "C" void vxref_normalize(const float* v, float* out) { float3 r = normalize(make_float3(v[0], v[1], v[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z; }
extern "C" void vxref_reflect(const float* i, const float* n, float* out) {
    float3 r = reflect(make_float3(i[0], i[1], i[2]), make_float3(n[0], n[1], n[2])); out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
extern "C" float vxref_dot(const float* a, const float* b) { return dot(make_float3(a[0], a[1], a[2]), make_float3(b[0], b[1], b[2])); }
extern "C" void vxref_export(vxref_world* w, uint32_t* top_words, uint32_t* brick_words, float* bounds6) {
    size_t wpb = (size_t)w->factor * w->factor * w->factor / 32;
    std::memcpy(top_words, w->top.grid.Raw(), (w->nbricks + 31) / 32 * 4);
    for (size_t i = 0; i < w->nbricks; ++i) {
        if (w->bricks[i].dimensions[0] == 0) std::memset(brick_words + i * wpb, 0, wpb * 4);
        else std::memcpy(brick_words + i * wpb, w->bricks[i].grid.Raw(), wpb * 4);
        const Bounds3Df& b = w->bounds[i];
        float* o = bounds6 + i * 6;
        o[0] = b.min.x; o[1] = b.min.y; o[2] = b.min.z; o[3] = b.max.x; o[4] = b.max.y; o[5] = b.max.z;
    }
}
extern "C" void vxref_trace(vxref_world* w, const float* o, const float* d, size_t n, int max_steps,
                            uint8_t* hit, float* pos, float* normal, int* steps, int nthreads) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    std::vector<std::thread> pool;
    for (int t = 0; t < nthreads; ++t)
        pool.emplace_back([=]() {
            for (size_t i = t; i < n; i += nthreads) {
                int s = 0; float3 nn = make_float3(0, 0, 0), pp = make_float3(0, 0, 0);
                bool h = Raytrace(max_steps, make_float3(o[3 * i], o[3 * i + 1], o[3 * i + 2]),
                                  make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]), w->top, w->bricks, w->bounds,
                                  w->factor, s, nn, pp);
                hit[i] = h; steps[i] = s;
                /* dispatch kernel semantics (VolumeRaytracer.cu:105-115) */
                pos[3 * i] = h ? pp.x : FLT_INF; pos[3 * i + 1] = h ? pp.y : FLT_INF; pos[3 * i + 2] = h ? pp.z : FLT_INF;
                normal[3 * i] = nn.x; normal[3 * i + 1] = nn.y; normal[3 * i + 2] = nn.z;
            }
        });
    for (auto& th : pool) th.join();
}
/* single-level DDARayTraversal over the dense grid, default params */
extern "C" void vxref_dda_dense(vxref_world* w, const float* s, const float* d, size_t n, uint8_t* hit, uint8_t* oob, int* steps) {
    for (size_t i = 0; i < n; ++i) {
        auto p = DDARayParams<float3, 3>::Default(w->dense, make_float3(s[3 * i], s[3 * i + 1], s[3 * i + 2]),
                                                  make_float3(d[3 * i], d[3 * i + 1], d[3 * i + 2]));
        DDARayResults<float3> r;
        DDARayTraversal(p, r);
        hit[i] = r.hit; oob[i] = r.isOutOfBounds; steps[i] = r.stepsTaken;
    }
}
extern "C" int vxref_ray_aabb(const float* s, const float* d, const float* lo, const float* hi, float* p, float* nrm) {
    float3 pp = make_float3(0, 0, 0), nn = make_float3(0, 0, 0);
    bool h = RayIntersectsAABB(make_float3(s[0], s[1], s[2]), make_float3(d[0], d[1], d[2]), make_float3(lo[0], lo[1], lo[2]),
                               make_float3(hi[0], hi[1], hi[2]), &pp, &nn);
    p[0] = pp.x; p[1] = pp.y; p[2] = pp.z; nrm[0] = nn.x; nrm[1] = nn.y; nrm[2] = nn.z;
    return h;
}
extern "C" uint32_t vxref_sample_index(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t h) { return GetSampleIndex(x, y, z, w, h); }
/* one DDARayTraversal call with explicit mode, for unit tests: level 0 = coarse (per-brick AABBs),
 * level 1 = fine inside brick `brick_index` (bounds [0,factor]^3) */
extern "C" void vxref_dda(vxref_world* w, int level, uint32_t brick_index, const float* s, const float* d,
                          int* flags, int* cells6, float* pn6, int* steps) {
    DDARayResults<float3> r;
    std::memset(&r, 0, sizeof(r));
    float3 S = make_float3(s[0], s[1], s[2]), D = make_float3(d[0], d[1], d[2]);
    Bounds3Df cb{};
    if (level == 0) {
        auto p = DDARayParams<float3, 3>::Default(w->top, S, D);
        p.per_voxel_bounds = w->bounds; p.per_voxel_bounds_scale = w->factor;
        DDARayTraversal(p, r);
    } else {
        auto p = DDARayParams<float3, 3>::Default(w->bricks[brick_index], S, D);
        cb.min = make_float3(0, 0, 0); cb.max = make_float3(w->factor, w->factor, w->factor);
        p.bounds = &cb;
        DDARayTraversal(p, r);
    }
    flags[0] = r.hit; flags[1] = r.isOutOfBounds; *steps = r.stepsTaken;
    cells6[0] = (int)r.HitCell.x; cells6[1] = (int)r.HitCell.y; cells6[2] = (int)r.HitCell.z;
    cells6[3] = (int)r.NextCell.x; cells6[4] = (int)r.NextCell.y; cells6[5] = (int)r.NextCell.z;
    pn6[0] = r.HitIntersectedPoint.x; pn6[1] = r.HitIntersectedPoint.y; pn6[2] = r.HitIntersectedPoint.z;
    pn6[3] = r.HitNormal.x; pn6[4] = r.HitNormal.y; pn6[5] = r.HitNormal.z;
}
