// vx_device.cuh — device-side brickmap view and the exact-arithmetic DDA primitives.
//
// Semantics follow the reference's VolumeRaytracer.cu:124-174 (slab test), :176-352 (one DDA
// level) and :354-525 (two-level march) under IEEE binary32 with no FMA contraction (SURVEY.md
// §8c).  Every product/sum that feeds a comparison or a float->int conversion goes through
// __fmul_rn/__fadd_rn/__fdiv_rn, so the results do not depend on -fmad or -use_fast_math.
// The memory layout is new: coarse occupancy bits, one 8-byte record per coarse
// cell (packed bounds + pool slot) and a compact pool of 64-byte-tile brick masks.
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>

namespace vx {

constexpr int kMaxSteps = 2048;        // VolumeRaytracer.cuh:235
constexpr float kFltEps = 1.1920929e-07f; // std::numeric_limits<float>::epsilon(), VolumeRaytracer.cuh:22
constexpr double kEpsDDA = 1e-6;       // VolumeRaytracer.cuh:20 (a double)

struct WorldView {
    const uint32_t* __restrict__ top;   // coarse occupancy bits, reference tiled index
    const uint2* __restrict__ records;  // per coarse cell: x = packed bounds (5 bits x 6), y = pool slot
    const uint32_t* __restrict__ pool;  // occupied bricks only, words_per_brick words each, in-brick tiled index
    int cx, cy, cz;                     // coarse dims
    int factor;
    float inv_factor;                   // 1 / factor when factor is a power of two (then x * inv_factor == x / factor exactly), else 0
    uint32_t words_per_brick;           // factor^3 / 32
    uint32_t tiles_x, tiles_xy;         // cx/8, (cx/8)*(cy/8)
    // traversal copies in our own layout: x-fastest linear grid with a one-cell border on every side
    // (dims cx+2, cy+2, cz+2); border bits are 1 so that leaving the grid shows up as an "occupied" lookup
    const uint32_t* __restrict__ ptop;  // padded occupancy bits
    const uint2* __restrict__ precords; // padded records (same packing as `records`)
    int pnx, pnxy;                      // cx+2, (cx+2)*(cy+2)
    uint32_t magic_x, shift_x;          // n / pnx  == __umulhi(n, magic_x)  >> shift_x   for n < 2^31
    uint32_t magic_xy, shift_xy;        // n / pnxy == __umulhi(n, magic_xy) >> shift_xy  for n < 2^31
};

struct Work { uint32_t C, K, E; };

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ float fmul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float fadd(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float fsub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fdiv(float a, float b) { return __fdiv_rn(a, b); }
__device__ __forceinline__ float dot3(V3 a, V3 b) { return fadd(fadd(fmul(a.x, b.x), fmul(a.y, b.y)), fmul(a.z, b.z)); }

template <int NORM>
__device__ __forceinline__ V3 normalize3(V3 v) {
    if (NORM == 2) return v;
    const float d = dot3(v, v);
    float inv;
    if (NORM == 1) inv = (float)__ddiv_rn(1.0, __dsqrt_rn((double)d)); // CUDA's HOST rsqrtf (crt/math_functions.hpp)
    else inv = rsqrtf(d);
    return mk(fmul(v.x, inv), fmul(v.y, inv), fmul(v.z, inv));
}

__device__ __forceinline__ uint32_t tiled_index(int x, int y, int z, uint32_t tiles_x, uint32_t tiles_xy) {
    // VolumeRaytracer.cuh:111-131 for non-negative coordinates
    const uint32_t tile = (uint32_t)(x >> 3) + (uint32_t)(y >> 3) * tiles_x + (uint32_t)(z >> 3) * tiles_xy;
    return tile * 512u + (uint32_t)(x & 7) + ((uint32_t)(y & 7) << 3) + ((uint32_t)(z & 7) << 6);
}

// ---------------------------------------------------------------------------------------------
// DDA primitives: branch-free step, per-ray invariants hoisted.  Same arithmetic, same order of
// float operations per value as the reference; only control flow and evaluation laziness differ
// (values the reference computes and never reads are not computed).  The march built from them
// is vx_march.cuh.
// ---------------------------------------------------------------------------------------------
struct Ray {
    bool dnz, dpos;        // every component of d finite and not zero / and positive (vx_march.cuh F_DNZ, F_DPOS)
    V3 d;
    float ivx, ivy, ivz;   // slab test: 1/(d == 0 ? FLT_EPS : d)          VolumeRaytracer.cu:127-129
    float dtx, dty, dtz;   // |1/d| or +inf                                :199-201
    int sx, sy, sz;        // d > 0 ? 1 : -1                               :195-197
};
__device__ __forceinline__ Ray make_ray(V3 d) {
    Ray r;
    r.dnz = false; r.dpos = false;
    r.d = d;
    r.ivx = fdiv(1.0f, d.x == 0.0f ? kFltEps : d.x);
    r.ivy = fdiv(1.0f, d.y == 0.0f ? kFltEps : d.y);
    r.ivz = fdiv(1.0f, d.z == 0.0f ? kFltEps : d.z);
    // for d != 0 the slab inverse IS 1/d, so |1/d| reuses it bit for bit
    r.dtx = d.x != 0.0f ? fabsf(r.ivx) : CUDART_INF_F;
    r.dty = d.y != 0.0f ? fabsf(r.ivy) : CUDART_INF_F;
    r.dtz = d.z != 0.0f ? fabsf(r.ivz) : CUDART_INF_F;
    r.sx = d.x > 0.0f ? 1 : -1; r.sy = d.y > 0.0f ? 1 : -1; r.sz = d.z > 0.0f ? 1 : -1;
    return r;
}

struct Slab { bool hit; V3 point; V3 normal; uint32_t ncode; };
__device__ __forceinline__ Slab ray_aabb(V3 s, const Ray& r, V3 lo, V3 hi) {
    const float ax = fmul(fsub(lo.x, s.x), r.ivx), bx = fmul(fsub(hi.x, s.x), r.ivx);
    const float ay = fmul(fsub(lo.y, s.y), r.ivy), by = fmul(fsub(hi.y, s.y), r.ivy);
    const float az = fmul(fsub(lo.z, s.z), r.ivz), bz = fmul(fsub(hi.z, s.z), r.ivz);
    const float nx = fminf(ax, bx), fx = fmaxf(ax, bx);
    const float ny = fminf(ay, by), fy = fmaxf(ay, by);
    const float nz = fminf(az, bz), fz = fmaxf(az, bz);
    const float t_in = fmaxf(fmaxf(nx, ny), nz);
    const float t_out = fminf(fminf(fx, fy), fz);
    Slab o;
    o.hit = !(t_out < fmaxf(t_in, 0.0f));
    o.point = mk(fadd(s.x, fmul(t_in, r.d.x)), fadd(s.y, fmul(t_in, r.d.y)), fadd(s.z, fmul(t_in, r.d.z)));
    if (t_in == nx) o.normal = mk(r.ivx < 0.0f ? -1.0f : 1.0f, 0.0f, 0.0f);
    else if (t_in == ny) o.normal = mk(0.0f, r.ivy < 0.0f ? -1.0f : 1.0f, 0.0f);
    else o.normal = mk(0.0f, 0.0f, r.ivz < 0.0f ? -1.0f : 1.0f);
    // the same normal as a 3-bit code (vx_march.cuh): 1 + 2*axis + (negative ? 1 : 0)
    o.ncode = t_in == nx ? (r.ivx < 0.0f ? 2u : 1u) : (t_in == ny ? (r.ivy < 0.0f ? 4u : 3u) : (r.ivz < 0.0f ? 6u : 5u));
    return o;
}

// Running state of one DDA level.
struct Dda {
    int cx, cy, cz;
    float tx, ty, tz;
    int limx, limy, limz; // exclusive upper cell limits incl. the reference's edge padding (:216-232)
};
__device__ __forceinline__ Dda dda_begin(V3 s, const Ray& r, int cols, int rows, int depth) {
    Dda q;
    q.cx = (int)s.x; q.cy = (int)s.y; q.cz = (int)s.z; // cvt.rzi: truncation, saturating
    q.tx = r.d.x != 0.0f ? fdiv(fsub((float)(q.cx + (r.sx > 0)), s.x), r.d.x) : CUDART_INF_F;
    q.ty = r.d.y != 0.0f ? fdiv(fsub((float)(q.cy + (r.sy > 0)), s.y), r.d.y) : CUDART_INF_F;
    q.tz = r.d.z != 0.0f ? fdiv(fsub((float)(q.cz + (r.sz > 0)), s.z), r.d.z) : CUDART_INF_F;
    const bool on_edge = q.cx == cols || q.cy == rows || q.cz == depth;
    q.limx = cols + ((on_edge && r.d.x < 0.0f) ? 1 : 0);
    q.limy = rows + ((on_edge && r.d.y < 0.0f) ? 1 : 0);
    q.limz = depth + ((on_edge && r.d.z < 0.0f) ? 1 : 0);
    return q;
}
__device__ __forceinline__ bool dda_inside(const Dda& q) {
    return (unsigned)q.cx < (unsigned)q.limx && (unsigned)q.cy < (unsigned)q.limy && (unsigned)q.cz < (unsigned)q.limz;
}
// One cell along the smallest tMax (ties: y over x, z over both — :293-322), without branches.
// Returns the axis masks and the parameter t of the crossing.
__device__ __forceinline__ void dda_step(Dda& q, const Ray& r, bool& mx, bool& my, float& t) {
    mx = q.tx < q.ty && q.tx < q.tz;
    my = !mx && q.ty <= q.tx && q.ty < q.tz;
    const bool mz = !mx && !my;
    t = mx ? q.tx : (my ? q.ty : q.tz);
    const float tn = fadd(t, mx ? r.dtx : (my ? r.dty : r.dtz));
    q.tx = mx ? tn : q.tx; q.ty = my ? tn : q.ty; q.tz = mz ? tn : q.tz;
    q.cx += mx ? r.sx : 0; q.cy += my ? r.sy : 0; q.cz += mz ? r.sz : 0;
}

} // namespace vx
