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
    // traversal copies in our own layout: x-fastest linear grid with a one-cell border on every side
    // (dims cx+2, cy+2, cz+2); border bits are 1 so that leaving the grid shows up as an "occupied" lookup
    const uint32_t* __restrict__ ptop;  // padded occupancy bits
    const uint2* __restrict__ precords; // padded records (same packing as `records`)
    int pnx, pnxy;                      // cx+2, (cx+2)*(cy+2)
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
// v2 DDA: branch-free step, per-ray invariants hoisted.  Same arithmetic, same
// order of float operations per value as the reference; only control flow and evaluation
// laziness differ (values the reference computes and never reads are not computed).
// ---------------------------------------------------------------------------------------------
struct Ray {
    V3 d;
    float ivx, ivy, ivz;   // slab test: 1/(d == 0 ? FLT_EPS : d)          VolumeRaytracer.cu:127-129
    float dtx, dty, dtz;   // |1/d| or +inf                                :199-201
    int sx, sy, sz;        // d > 0 ? 1 : -1                               :195-197
};
__device__ __forceinline__ Ray make_ray(V3 d) {
    Ray r;
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

struct Slab { bool hit; V3 point; V3 normal; };
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

struct CoarseOut {
    bool hit, oob;
    int steps;
    int cell[3];
    V3 point, normal;
    uint2 rec;
    Dda at_exit; // state before the reference's final advance; NextCell = one more dda_step (rarely needed)
};

// Occupied coarse cell (qx,qy,qz): build the brick's tight AABB from its packed bounds and slab-test it
// ((min + 0) / scale + cell, (max + 1) / scale + cell — VolumeRaytracer.cu:248-273).
__device__ __forceinline__ bool coarse_candidate(const WorldView& w, V3 s, const Ray& r, int qx, int qy, int qz, uint2 rec, int it, CoarseOut& o) {
    const uint32_t b = rec.x;
    const float l0 = (float)(b & 31u), l1 = (float)((b >> 5) & 31u), l2 = (float)((b >> 10) & 31u);
    const float h0 = (float)(((b >> 15) & 31u) + 1u), h1 = (float)(((b >> 20) & 31u) + 1u), h2 = (float)(((b >> 25) & 31u) + 1u);
    V3 lo, hi;
    if ((w.factor & (w.factor - 1)) == 0) { // x / 2^k == x * 2^-k exactly
        const float inv_f = fdiv(1.0f, (float)w.factor);
        lo = mk(fadd(fmul(l0, inv_f), (float)qx), fadd(fmul(l1, inv_f), (float)qy), fadd(fmul(l2, inv_f), (float)qz));
        hi = mk(fadd(fmul(h0, inv_f), (float)qx), fadd(fmul(h1, inv_f), (float)qy), fadd(fmul(h2, inv_f), (float)qz));
    } else {
        const float fs = (float)w.factor;
        lo = mk(fadd(fdiv(l0, fs), (float)qx), fadd(fdiv(l1, fs), (float)qy), fadd(fdiv(l2, fs), (float)qz));
        hi = mk(fadd(fdiv(h0, fs), (float)qx), fadd(fdiv(h1, fs), (float)qy), fadd(fdiv(h2, fs), (float)qz));
    }
    const Slab sl = ray_aabb(s, r, lo, hi);
    if (!sl.hit) return false;
    o.hit = true; o.normal = sl.normal;
    if (it != 0) o.point = sl.point;
    o.cell[0] = qx; o.cell[1] = qy; o.cell[2] = qz;
    o.rec = rec;
    return true;
}

// Coarse level: cells are bricks.  The crossing point / normal of plain steps are never read by Raytrace
// for the coarse level (only a hit's AABB point/normal are), so they are not formed.
//
// Two loops with identical step semantics:
//  * a generic one (explicit bounds test, clamped lookups) that runs while the current cell is not strictly
//    inside the grid — i.e. the ray starts outside, or sits on the reference's padded far face (:216-232);
//  * the hot one over the bordered linear grid: the cell index advances by a per-axis constant, leaving the
//    grid is detected by the border's sentinel bits, and cell coordinates are only recovered at events.
template <bool COUNT>
__device__ __forceinline__ void dda_coarse(const WorldView& w, V3 s, const Ray& r, CoarseOut& o, Work& wk) {
    Dda q = dda_begin(s, r, w.cx, w.cy, w.cz);
    o.hit = false; o.oob = false; o.steps = 0;
    o.point = s; o.normal = mk(0.0f, 0.0f, 0.0f);
    o.cell[0] = o.cell[1] = o.cell[2] = 0;
    o.rec = make_uint2(0u, 0u);
    int it = 0;
    bool done = false;
    while (it < kMaxSteps && !((unsigned)q.cx < (unsigned)w.cx && (unsigned)q.cy < (unsigned)w.cy && (unsigned)q.cz < (unsigned)w.cz)) {
        if (!dda_inside(q)) { o.oob = true; done = true; break; }
        const int qx = min(q.cx, w.cx - 1), qy = min(q.cy, w.cy - 1), qz = min(q.cz, w.cz - 1);
        const uint32_t idx = (uint32_t)((qz + 1) * w.pnxy + (qy + 1) * w.pnx + (qx + 1));
        const bool occ = (__ldg(w.ptop + (idx >> 5)) >> (idx & 31u)) & 1u;
        if (COUNT) { wk.C++; wk.K += occ; }
        if (occ && coarse_candidate(w, s, r, qx, qy, qz, __ldg(w.precords + idx), it, o)) { done = true; break; }
        bool mx, my; float t;
        dda_step(q, r, mx, my, t);
        ++it;
    }
    if (!done && it < kMaxSteps) {
        uint32_t idx = (uint32_t)((q.cz + 1) * w.pnxy + (q.cy + 1) * w.pnx + (q.cx + 1));
        const int dix = r.sx, diy = r.sy * w.pnx, diz = r.sz * w.pnxy;
        float tx = q.tx, ty = q.ty, tz = q.tz;
        for (; it < kMaxSteps; ++it) {
            if ((__ldg(w.ptop + (idx >> 5)) >> (idx & 31u)) & 1u) {
                const int pz = (int)(idx / (uint32_t)w.pnxy), rem = (int)(idx - (uint32_t)pz * (uint32_t)w.pnxy);
                const int py = rem / w.pnx, px = rem - py * w.pnx;
                q.cx = px - 1; q.cy = py - 1; q.cz = pz - 1;
                if (!((unsigned)q.cx < (unsigned)w.cx && (unsigned)q.cy < (unsigned)w.cy && (unsigned)q.cz < (unsigned)w.cz)) { o.oob = true; break; }
                if (COUNT) { wk.C++; wk.K++; }
                if (coarse_candidate(w, s, r, q.cx, q.cy, q.cz, __ldg(w.precords + idx), it, o)) break;
            } else if (COUNT) wk.C++;
            const bool mx = tx < ty && tx < tz;
            const bool my = !mx && ty <= tx && ty < tz;
            const bool mz = !mx && !my;
            const float t = mx ? tx : (my ? ty : tz);
            const float tn = fadd(t, mx ? r.dtx : (my ? r.dty : r.dtz));
            tx = mx ? tn : tx; ty = my ? tn : ty; tz = mz ? tn : tz;
            idx += (uint32_t)(mx ? dix : (my ? diy : diz));
        }
        q.tx = tx; q.ty = ty; q.tz = tz;
    }
    o.steps = it;
    o.at_exit = q; // cell + tMax before the reference's final advance (cell valid whenever hit or oob)
}

struct FineOut {
    bool hit, oob;
    int steps;
    int cell[3];
    V3 point, normal;
};

// Fine level: cells are voxels of one brick; crossing points are tested against [0,factor]^3 before a step
// is counted (VolumeRaytracer.cu:325-341).  FINE8: factor == 8, mask bit = x + 8y + 64z.
template <bool FINE8>
__device__ __forceinline__ void dda_fine(const uint32_t* __restrict__ mask, int f, V3 s, const Ray& r, FineOut& o) {
    Dda q = dda_begin(s, r, f, f, f);
    o.hit = false; o.oob = false;
    o.point = s; o.normal = mk(0.0f, 0.0f, 0.0f);
    o.cell[0] = o.cell[1] = o.cell[2] = 0;
    const uint32_t tiles = (uint32_t)f >> 3;
    const float hi = (float)f;
    const int ox = r.sx > 0, oy = r.sy > 0, oz = r.sz > 0;
    int steps = 0;
    for (int it = 0; it < kMaxSteps; ++it) {
        if (!dda_inside(q)) { o.oob = true; break; }
        const int qx = min(q.cx, f - 1), qy = min(q.cy, f - 1), qz = min(q.cz, f - 1);
        const uint32_t idx = FINE8 ? (uint32_t)(qx + (qy << 3) + (qz << 6)) : tiled_index(qx, qy, qz, tiles, tiles * tiles);
        if ((__ldg(mask + (idx >> 5)) >> (idx & 31u)) & 1u) {
            o.hit = true; o.cell[0] = qx; o.cell[1] = qy; o.cell[2] = qz;
            break;
        }
        const float fx = (float)(q.cx + ox), fy = (float)(q.cy + oy), fz = (float)(q.cz + oz); // faces of the cell being left
        bool mx, my; float t;
        dda_step(q, r, mx, my, t);
        const bool mz = !mx && !my;
        const V3 p = mk(mx ? fx : fadd(s.x, fmul(t, r.d.x)), my ? fy : fadd(s.y, fmul(t, r.d.y)), mz ? fz : fadd(s.z, fmul(t, r.d.z)));
        if (p.x < 0.0f || p.x > hi || p.y < 0.0f || p.y > hi || p.z < 0.0f || p.z > hi) { o.oob = true; break; }
        ++steps;
        o.point = p;
        o.normal = mk(mx ? (float)r.sx : 0.0f, my ? (float)r.sy : 0.0f, mz ? (float)r.sz : 0.0f);
    }
    o.steps = steps;
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
    const Ray r = make_ray(normalize3<NORM>(ray));
    V3 start_normal = mk(0.0f, 0.0f, 0.0f);
    if (!(start.x >= 0.0f && start.y >= 0.0f && start.z >= 0.0f && start.x < (float)w.cx && start.y < (float)w.cy &&
          start.z < (float)w.cz)) {
        const float e = (float)kEpsDDA;
        const Slab sl = ray_aabb(start, r, mk(e, e, e),
                                 mk((float)((double)w.cx - kEpsDDA), (float)((double)w.cy - kEpsDDA), (float)((double)w.cz - kEpsDDA)));
        if (sl.hit) { start = sl.point; start_normal = sl.normal; }
    }
    V3 normal = mk(0.0f, 0.0f, 0.0f), hit_pos = mk(0.0f, 0.0f, 0.0f);
    bool hit = false;
    while (total < max_steps) {
        CoarseOut c;
        dda_coarse<COUNT>(w, start, r, c, o.work);
        total += c.steps; o.coarse_steps += c.steps;
        if (!c.hit) break; // miss or out of bounds (:508-511); hitPosition is not read on a miss
        if (pcx == c.cell[0] && pcy == c.cell[1] && pcz == c.cell[2]) break; // :402-407
        pcx = c.cell[0]; pcy = c.cell[1]; pcz = c.cell[2];
        const float bx = fmul((float)pcx, ff), by = fmul((float)pcy, ff), bz = fmul((float)pcz, ff);
        const V3 local = mk(fsub(fmul(c.point.x, ff), bx), fsub(fmul(c.point.y, ff), by), fsub(fmul(c.point.z, ff), bz));
        if (COUNT) o.work.E++;
        FineOut q;
        const uint32_t* mask = w.pool + (size_t)c.rec.y * w.words_per_brick;
        if (w.factor == 8) dda_fine<true>(mask, 8, local, r, q);
        else dda_fine<false>(mask, w.factor, local, r, q);
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
        if (q.oob) { // :438-489
            if (pcx == (int)start.x && pcy == (int)start.y && pcz == (int)start.z) {
                start.x = nextafterf(start.x, r.d.x < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                start.y = nextafterf(start.y, r.d.y < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                start.z = nextafterf(start.z, r.d.z < 0.0f ? -CUDART_INF_F : CUDART_INF_F);
                if (pcx == (int)start.x && pcy == (int)start.y && pcz == (int)start.z) {
                    Dda n = c.at_exit; // NextCell of the COARSE traversal = the cell after its final advance (:347,:473)
                    bool mx, my; float t;
                    dda_step(n, r, mx, my, t);
                    const float gx = fsub((float)n.cx, start.x), gy = fsub((float)n.cy, start.y), gz = fsub((float)n.cz, start.z);
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
