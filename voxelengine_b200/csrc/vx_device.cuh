// vx_device.cuh — device-side brickmap view and the exact-arithmetic two-level DDA.
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
    uint32_t words_per_brick;           // factor^3 / 32
    uint32_t tiles_x, tiles_xy;         // cx/8, (cx/8)*(cy/8)
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

struct Slab { bool hit; V3 point; V3 normal; };
__device__ __forceinline__ Slab ray_aabb(V3 s, V3 d, V3 lo, V3 hi) {
    const float ivx = fdiv(1.0f, d.x == 0.0f ? kFltEps : d.x);
    const float ivy = fdiv(1.0f, d.y == 0.0f ? kFltEps : d.y);
    const float ivz = fdiv(1.0f, d.z == 0.0f ? kFltEps : d.z);
    const float ax = fmul(fsub(lo.x, s.x), ivx), bx = fmul(fsub(hi.x, s.x), ivx);
    const float ay = fmul(fsub(lo.y, s.y), ivy), by = fmul(fsub(hi.y, s.y), ivy);
    const float az = fmul(fsub(lo.z, s.z), ivz), bz = fmul(fsub(hi.z, s.z), ivz);
    const float nx = fminf(ax, bx), fx = fmaxf(ax, bx);
    const float ny = fminf(ay, by), fy = fmaxf(ay, by);
    const float nz = fminf(az, bz), fz = fmaxf(az, bz);
    const float t_in = fmaxf(fmaxf(nx, ny), nz);
    const float t_out = fminf(fminf(fx, fy), fz);
    Slab r;
    r.hit = !(t_out < fmaxf(t_in, 0.0f));
    r.point = mk(fadd(s.x, fmul(t_in, d.x)), fadd(s.y, fmul(t_in, d.y)), fadd(s.z, fmul(t_in, d.z)));
    if (t_in == nx) r.normal = mk(ivx < 0.0f ? -1.0f : 1.0f, 0.0f, 0.0f);
    else if (t_in == ny) r.normal = mk(0.0f, ivy < 0.0f ? -1.0f : 1.0f, 0.0f);
    else r.normal = mk(0.0f, 0.0f, ivz < 0.0f ? -1.0f : 1.0f);
    return r;
}

struct DdaOut {
    bool hit, oob;
    int cell[3];
    int next[3];
    V3 point, normal;
    int steps;
};

// Per-ray constants of one DDA level.
struct DdaSetup {
    int cx, cy, cz, sx, sy, sz;
    float tx, ty, tz, dtx, dty, dtz;
    int limx, limy, limz; // exclusive upper cell limits incl. the reference's edge padding
};
__device__ __forceinline__ DdaSetup dda_setup(V3 s, V3 d, int cols, int rows, int depth) {
    DdaSetup q;
    q.cx = (int)s.x; q.cy = (int)s.y; q.cz = (int)s.z; // cvt.rzi: truncation, saturating
    q.sx = d.x > 0.0f ? 1 : -1; q.sy = d.y > 0.0f ? 1 : -1; q.sz = d.z > 0.0f ? 1 : -1;
    q.dtx = d.x != 0.0f ? fabsf(fdiv(1.0f, d.x)) : CUDART_INF_F;
    q.dty = d.y != 0.0f ? fabsf(fdiv(1.0f, d.y)) : CUDART_INF_F;
    q.dtz = d.z != 0.0f ? fabsf(fdiv(1.0f, d.z)) : CUDART_INF_F;
    q.tx = d.x != 0.0f ? fdiv(fsub((float)(q.cx + (q.sx > 0)), s.x), d.x) : CUDART_INF_F;
    q.ty = d.y != 0.0f ? fdiv(fsub((float)(q.cy + (q.sy > 0)), s.y), d.y) : CUDART_INF_F;
    q.tz = d.z != 0.0f ? fdiv(fsub((float)(q.cz + (q.sz > 0)), s.z), d.z) : CUDART_INF_F;
    const bool on_edge = q.cx == cols || q.cy == rows || q.cz == depth; // VolumeRaytracer.cu:216-232
    q.limx = cols + ((on_edge && d.x < 0.0f) ? 1 : 0);
    q.limy = rows + ((on_edge && d.y < 0.0f) ? 1 : 0);
    q.limz = depth + ((on_edge && d.z < 0.0f) ? 1 : 0);
    return q;
}

// Advance one cell along the smallest tMax (ties: y over x, z over both — VolumeRaytracer.cu:293-322).
// Returns the axis advanced (0,1,2) and the face-crossing point.
__device__ __forceinline__ int dda_advance(DdaSetup& q, V3 s, V3 d, V3& p) {
    if (q.tx < q.ty && q.tx < q.tz) {
        p = mk((float)(q.cx + (q.sx > 0)), fadd(s.y, fmul(q.tx, d.y)), fadd(s.z, fmul(q.tx, d.z)));
        q.cx += q.sx; q.tx = fadd(q.tx, q.dtx);
        return 0;
    } else if (q.ty <= q.tx && q.ty < q.tz) {
        p = mk(fadd(s.x, fmul(q.ty, d.x)), (float)(q.cy + (q.sy > 0)), fadd(s.z, fmul(q.ty, d.z)));
        q.cy += q.sy; q.ty = fadd(q.ty, q.dty);
        return 1;
    } else {
        p = mk(fadd(s.x, fmul(q.tz, d.x)), fadd(s.y, fmul(q.tz, d.y)), (float)(q.cz + (q.sz > 0)));
        q.cz += q.sz; q.tz = fadd(q.tz, q.dtz);
        return 2;
    }
}
__device__ __forceinline__ V3 axis_normal(int axis, const DdaSetup& q) {
    return axis == 0 ? mk((float)q.sx, 0.0f, 0.0f) : (axis == 1 ? mk(0.0f, (float)q.sy, 0.0f) : mk(0.0f, 0.0f, (float)q.sz));
}

// Coarse level: cells are bricks; a cell is a candidate when its occupancy bit is set and the ray
// hits the brick's tight AABB (VolumeRaytracer.cu:248-273).  `rec_out` returns the record of the hit cell.
template <bool COUNT>
__device__ __forceinline__ void dda_coarse(const WorldView& w, V3 s, V3 d, DdaOut& r, uint2& rec_out, Work& wk) {
    DdaSetup q = dda_setup(s, d, w.cx, w.cy, w.cz);
    r.point = s; r.hit = false; r.oob = false; r.steps = 0;
    r.cell[0] = r.cell[1] = r.cell[2] = 0;
    r.normal = mk(0.0f, 0.0f, 0.0f);
    const float fs = (float)w.factor;
    bool leave = false;
    for (int it = 0; it < kMaxSteps; ++it) {
        if (0 <= q.cx && q.cx < q.limx && 0 <= q.cy && q.cy < q.limy && 0 <= q.cz && q.cz < q.limz) {
            const int qx = min(q.cx, w.cx - 1), qy = min(q.cy, w.cy - 1), qz = min(q.cz, w.cz - 1);
            r.cell[0] = qx; r.cell[1] = qy; r.cell[2] = qz;
            const uint32_t idx = tiled_index(qx, qy, qz, w.tiles_x, w.tiles_xy);
            const bool occ = (__ldg(w.top + (idx >> 5)) >> (idx & 31u)) & 1u;
            if (COUNT) { wk.C++; wk.K += occ; }
            if (occ) {
                const uint2 rec = __ldg(w.records + idx);
                const uint32_t b = rec.x;
                const V3 lo = mk(fadd(fdiv((float)(b & 31u), fs), (float)qx), fadd(fdiv((float)((b >> 5) & 31u), fs), (float)qy),
                                 fadd(fdiv((float)((b >> 10) & 31u), fs), (float)qz));
                const V3 hi = mk(fadd(fdiv((float)(((b >> 15) & 31u) + 1u), fs), (float)qx),
                                 fadd(fdiv((float)(((b >> 20) & 31u) + 1u), fs), (float)qy),
                                 fadd(fdiv((float)(((b >> 25) & 31u) + 1u), fs), (float)qz));
                const Slab sl = ray_aabb(s, d, lo, hi);
                if (sl.hit) {
                    r.hit = true; r.normal = sl.normal;
                    if (it != 0) r.point = sl.point;
                    rec_out = rec;
                    leave = true;
                }
            }
        } else {
            r.oob = true; leave = true;
        }
        V3 p;
        const int axis = dda_advance(q, s, d, p);
        if (leave) break;
        r.normal = axis_normal(axis, q);
        r.steps += 1;
        r.point = p;
    }
    r.next[0] = q.cx; r.next[1] = q.cy; r.next[2] = q.cz;
}

// Fine level: cells are voxels of one brick (factor^3 bits, in-brick tiled index); crossing points
// are tested against [0,factor]^3 before a step is counted (VolumeRaytracer.cu:325-341).
__device__ __forceinline__ void dda_fine(const uint32_t* __restrict__ mask, int f, V3 s, V3 d, DdaOut& r) {
    DdaSetup q = dda_setup(s, d, f, f, f);
    r.point = s; r.hit = false; r.oob = false; r.steps = 0;
    r.cell[0] = r.cell[1] = r.cell[2] = 0;
    r.normal = mk(0.0f, 0.0f, 0.0f);
    const uint32_t tiles = (uint32_t)f >> 3;
    const float hi = (float)f;
    bool leave = false;
    for (int it = 0; it < kMaxSteps; ++it) {
        if (0 <= q.cx && q.cx < q.limx && 0 <= q.cy && q.cy < q.limy && 0 <= q.cz && q.cz < q.limz) {
            const int qx = min(q.cx, f - 1), qy = min(q.cy, f - 1), qz = min(q.cz, f - 1);
            r.cell[0] = qx; r.cell[1] = qy; r.cell[2] = qz;
            const uint32_t idx = tiled_index(qx, qy, qz, tiles, tiles * tiles);
            if ((__ldg(mask + (idx >> 5)) >> (idx & 31u)) & 1u) { r.hit = true; leave = true; }
        } else {
            r.oob = true; leave = true;
        }
        V3 p;
        const int axis = dda_advance(q, s, d, p);
        if (leave) break;
        r.normal = axis_normal(axis, q);
        if (p.x < 0.0f || p.x > hi || p.y < 0.0f || p.y > hi || p.z < 0.0f || p.z > hi) { r.oob = true; break; }
        r.steps += 1;
        r.point = p;
    }
    r.next[0] = q.cx; r.next[1] = q.cy; r.next[2] = q.cz;
}

struct MarchOut {
    bool hit;
    int steps, coarse_steps, fine_steps;
    V3 pos, normal;
    int voxel[3];
    uint32_t brick;
    Work work;
};

// Two-level march, VolumeRaytracer.cu:354-525.
template <int NORM, bool COUNT>
__device__ __forceinline__ bool march(const WorldView& w, V3 origin, V3 ray, int max_steps, MarchOut& o) {
    const float ff = (float)w.factor;
    int pcx = -1, pcy = -1, pcz = -1;
    int total = 0;
    o.hit = false; o.coarse_steps = 0; o.fine_steps = 0; o.brick = 0xFFFFFFFFu;
    o.voxel[0] = o.voxel[1] = o.voxel[2] = -1;
    o.work.C = o.work.K = o.work.E = 0;
    V3 start = mk(fdiv(origin.x, ff), fdiv(origin.y, ff), fdiv(origin.z, ff));
    const V3 d = normalize3<NORM>(ray);
    V3 start_normal = mk(0.0f, 0.0f, 0.0f);
    if (!(start.x >= 0.0f && start.y >= 0.0f && start.z >= 0.0f && start.x < (float)w.cx && start.y < (float)w.cy &&
          start.z < (float)w.cz)) {
        const float e = (float)kEpsDDA;
        const Slab sl = ray_aabb(start, d, mk(e, e, e),
                                 mk((float)((double)w.cx - kEpsDDA), (float)((double)w.cy - kEpsDDA), (float)((double)w.cz - kEpsDDA)));
        if (sl.hit) { start = sl.point; start_normal = sl.normal; }
    }
    V3 normal = mk(0.0f, 0.0f, 0.0f), hit_pos = mk(0.0f, 0.0f, 0.0f);
    bool hit = false;
    while (total < max_steps) {
        DdaOut c;
        uint2 rec = make_uint2(0u, 0u);
        dda_coarse<COUNT>(w, start, d, c, rec, o.work);
        total += c.steps; o.coarse_steps += c.steps;
        V3 local = mk(fmul(c.point.x, ff), fmul(c.point.y, ff), fmul(c.point.z, ff));
        hit_pos = local;
        if (!(c.hit && !c.oob)) break;
        if (pcx == c.cell[0] && pcy == c.cell[1] && pcz == c.cell[2]) break; // VolumeRaytracer.cu:402-407
        pcx = c.cell[0]; pcy = c.cell[1]; pcz = c.cell[2];
        const float bx = fmul((float)pcx, ff), by = fmul((float)pcy, ff), bz = fmul((float)pcz, ff);
        local = mk(fsub(local.x, bx), fsub(local.y, by), fsub(local.z, bz));
        if (COUNT) o.work.E++;
        DdaOut q;
        dda_fine(w.pool + (size_t)rec.y * w.words_per_brick, w.factor, local, d, q);
        total += q.steps; o.fine_steps += q.steps;
        hit_pos = mk(fadd(q.point.x, bx), fadd(q.point.y, by), fadd(q.point.z, bz));
        if (q.hit) {
            normal = q.steps == 0 ? c.normal : q.normal;
            hit = true;
            o.brick = tiled_index(pcx, pcy, pcz, w.tiles_x, w.tiles_xy);
            o.voxel[0] = pcx * w.factor + q.cell[0]; o.voxel[1] = pcy * w.factor + q.cell[1]; o.voxel[2] = pcz * w.factor + q.cell[2];
            break;
        }
        start = mk(fdiv(hit_pos.x, ff), fdiv(hit_pos.y, ff), fdiv(hit_pos.z, ff));
        if (q.oob) { // VolumeRaytracer.cu:438-489
            const bool ex = pcx == (int)start.x, ey = pcy == (int)start.y, ez = pcz == (int)start.z;
            if (ex && ey && ez) {
                start.x = nextafterf(start.x, d.x < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                start.y = nextafterf(start.y, d.y < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                start.z = nextafterf(start.z, d.z < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                if (pcx == (int)start.x && pcy == (int)start.y && pcz == (int)start.z) {
                    const float gx = fsub((float)c.next[0], start.x), gy = fsub((float)c.next[1], start.y), gz = fsub((float)c.next[2], start.z);
                    const float ax = fabsf(gx), ay = fabsf(gy), az = fabsf(gz);
                    if (ax < ay && ax < az) start.x = fadd(start.x, gx);
                    else if (ay < ax && ay < az) start.y = fadd(start.y, gy);
                    else start.z = fadd(start.z, gz);
                }
            }
        }
    }
    o.steps = total;
    o.normal = normal;
    o.hit = hit;
    o.pos = hit_pos;
    if (hit && total == 0) { o.pos = mk(fmul(start.x, ff), fmul(start.y, ff), fmul(start.z, ff)); o.normal = start_normal; }
    return hit;
}

} // namespace vx
