/*
 * vxo_trace.cpp — CPU ORACLE (test infrastructure, not product code): slab test, one-level DDA,
 * two-level brickmap march, batch entry.
 *
 * Restates VolumeRaytracer.cu:124-174 (RayIntersectsAABB), :176-352 (DDARayTraversal),
 * :354-525 (Raytrace), :95-117 (dispatch) for the CPU, binary32 with no contraction, and in
 * addition returns the locals the parity contract compares: coarse cell, brick index (:419),
 * fine cell, per-level step counts, and the work counters C/K/E of SURVEY.md §8(d).
 * Parity status: pinned against the reference's own VolumeRaytracer.cu compiled for the host
 * (oracle/_ref/libvxref_host.so; tests/test_oracle.py, tests/test_reference_pins.py).
 */
#include "vxo.h"
#include "vxo_common.h"
#include "vxo_trace.h"
#include "vxo_world.h"

namespace vxo {

bool ray_aabb(V3 s, V3 d, V3 bmin, V3 bmax, V3* point, V3* normal) {
    const float ivx = 1.0f / (d.x == 0 ? kFltEps : d.x);
    const float ivy = 1.0f / (d.y == 0 ? kFltEps : d.y);
    const float ivz = 1.0f / (d.z == 0 ? kFltEps : d.z);
    const float ax = (bmin.x - s.x) * ivx, bx = (bmax.x - s.x) * ivx;
    const float ay = (bmin.y - s.y) * ivy, by = (bmax.y - s.y) * ivy;
    const float az = (bmin.z - s.z) * ivz, bz = (bmax.z - s.z) * ivz;
    const float nx = fmin_cu(ax, bx), fx = fmax_cu(ax, bx);
    const float ny = fmin_cu(ay, by), fy = fmax_cu(ay, by);
    const float nz = fmin_cu(az, bz), fz = fmax_cu(az, bz);
    const float t_in = fmax_cu(fmax_cu(nx, ny), nz);
    const float t_out = fmin_cu(fmin_cu(fx, fy), fz);
    if (t_out < fmax_cu(t_in, 0.0f)) return false;
    if (point) {
        float mx = t_in * d.x, my = t_in * d.y, mz = t_in * d.z;
        *point = V3{s.x + mx, s.y + my, s.z + mz};
    }
    if (normal) {
        if (t_in == nx) *normal = V3{ivx < 0.0f ? -1.0f : 1.0f, 0, 0};
        else if (t_in == ny) *normal = V3{0, ivy < 0.0f ? -1.0f : 1.0f, 0};
        else *normal = V3{0, 0, ivz < 0.0f ? -1.0f : 1.0f};
    }
    return true;
}

/* One level of Amanatides-Woo DDA.  `bounds6` != null selects the coarse mode (per-cell AABBs,
 * scale = factor); `clip_max` >= 0 selects the fine mode's crossing-point test against [0,clip_max]. */
void dda(const Level& L, V3 s, V3 d, const float* bounds6, int scale, int clip_max, DdaOut& r, Work* wk) {
    int cx = f2i(s.x), cy = f2i(s.y), cz = f2i(s.z);
    const int cols = L.nx, rows = L.ny, depth = L.nz;
    const int sx = d.x > 0 ? 1 : -1, sy = d.y > 0 ? 1 : -1, sz = d.z > 0 ? 1 : -1;
    const float dtx = d.x != 0 ? std::fabs(1.0f / d.x) : kInf;
    const float dty = d.y != 0 ? std::fabs(1.0f / d.y) : kInf;
    const float dtz = d.z != 0 ? std::fabs(1.0f / d.z) : kInf;
    float tx = d.x != 0 ? ((float)(cx + (sx > 0)) - s.x) / d.x : kInf;
    float ty = d.y != 0 ? ((float)(cy + (sy > 0)) - s.y) / d.y : kInf;
    float tz = d.z != 0 ? ((float)(cz + (sz > 0)) - s.z) / d.z : kInf;

    r.point = s; r.hit = false; r.oob = false; r.steps = 0;
    r.cell[0] = r.cell[1] = r.cell[2] = 0; r.next[0] = r.next[1] = r.next[2] = 0;
    r.normal = V3{0, 0, 0};
    bool leave = false;
    const bool on_edge = cx == cols || cy == rows || cz == depth;
    const float padx = (on_edge && d.x < 0) ? 1.0f : 0.0f;
    const float pady = (on_edge && d.y < 0) ? 1.0f : 0.0f;
    const float padz = (on_edge && d.z < 0) ? 1.0f : 0.0f;

    for (int it = 0; it < kMaxSteps; ++it) {
        if (0 <= cx && (float)cx < (float)cols + padx && 0 <= cy && (float)cy < (float)rows + pady &&
            0 <= cz && (float)cz < (float)depth + padz) {
            const int qx = std::min(std::max(cx, 0), cols - 1);
            const int qy = std::min(std::max(cy, 0), rows - 1);
            const int qz = std::min(std::max(cz, 0), depth - 1);
            r.cell[0] = qx; r.cell[1] = qy; r.cell[2] = qz;
            const uint64_t idx = sample_index((uint32_t)qx, (uint32_t)qy, (uint32_t)qz, (uint32_t)cols, (uint32_t)rows);
            const bool occ = bit_get(L.bits, idx);
            if (wk) { if (bounds6) { wk->C++; if (occ) wk->K++; } }
            if (bounds6) {
                const float* b = bounds6 + idx * 6;
                const float fs = (float)scale;
                const V3 lo{(b[0] + 0.0f) / fs + (float)qx, (b[1] + 0.0f) / fs + (float)qy, (b[2] + 0.0f) / fs + (float)qz};
                const V3 hi{(b[3] + 1.0f) / fs + (float)qx, (b[4] + 1.0f) / fs + (float)qy, (b[5] + 1.0f) / fs + (float)qz};
                if (occ && lo.x <= hi.x) {
                    V3 p, n;
                    if (ray_aabb(s, d, lo, hi, &p, &n)) {
                        r.hit = true; r.normal = n;
                        if (it != 0) r.point = p;
                        leave = true;
                    }
                }
            } else if (occ) {
                r.hit = true; leave = true;
            }
        } else {
            r.oob = true; leave = true;
        }

        float px, py, pz;
        if (tx < ty && tx < tz) {
            px = (float)(cx + (sx > 0)); py = s.y + tx * d.y; pz = s.z + tx * d.z;
            cx += sx; tx += dtx;
            if (!leave) r.normal = V3{(float)sx, 0, 0};
        } else if (ty <= tx && ty < tz) {
            px = s.x + ty * d.x; py = (float)(cy + (sy > 0)); pz = s.z + ty * d.z;
            cy += sy; ty += dty;
            if (!leave) r.normal = V3{0, (float)sy, 0};
        } else {
            px = s.x + tz * d.x; py = s.y + tz * d.y; pz = (float)(cz + (sz > 0));
            cz += sz; tz += dtz;
            if (!leave) r.normal = V3{0, 0, (float)sz};
        }
        if (leave) { r.next[0] = cx; r.next[1] = cy; r.next[2] = cz; break; }
        if (clip_max >= 0) {
            const float lo = 0.0f, hi = (float)clip_max;
            if (px < lo || px > hi || py < lo || py > hi || pz < lo || pz > hi) { r.oob = true; break; }
        }
        r.steps += 1;
        r.point = V3{px, py, pz};
    }
}

bool march(const vxo_world& w, V3 origin, V3 ray, int norm_mode, int max_steps, MarchOut& o) {
    const int f = w.factor;
    const float ff = (float)f;
    Level top{w.top_p, (int)w.cx, (int)w.cy, (int)w.cz};
    float prev[3] = {-1, -1, -1};
    int total = 0;
    o = MarchOut{};
    V3 start{origin.x / ff, origin.y / ff, origin.z / ff};
    const V3 d = norm_mode == VXO_NORM_NONE ? ray : normalize_host(ray);
    V3 start_normal{0, 0, 0};
    if (!(start.x >= 0 && start.y >= 0 && start.z >= 0 && start.x < (float)top.nx && start.y < (float)top.ny &&
          start.z < (float)top.nz)) {
        const float e = (float)kEpsDDA;
        V3 p;
        if (ray_aabb(start, d, V3{e, e, e},
                     V3{(float)((double)top.nx - kEpsDDA), (float)((double)top.ny - kEpsDDA), (float)((double)top.nz - kEpsDDA)},
                     &p, &start_normal))
            start = p;
    }
    V3 normal{0, 0, 0}, hit_pos{0, 0, 0};
    bool hit = false;
    while (total < max_steps) {
        DdaOut c;
        dda(top, start, d, w.bounds_p, f, -1, c, &o.work);
        total += c.steps; o.coarse_steps += c.steps;
        V3 local{c.point.x * ff, c.point.y * ff, c.point.z * ff};
        hit_pos = local;
        if (!(c.hit && !c.oob)) break;
        const float hc[3] = {(float)c.cell[0], (float)c.cell[1], (float)c.cell[2]};
        if (prev[0] == hc[0] && prev[1] == hc[1] && prev[2] == hc[2]) break;
        prev[0] = hc[0]; prev[1] = hc[1]; prev[2] = hc[2];
        const float bx = hc[0] * ff, by = hc[1] * ff, bz = hc[2] * ff;
        local.x -= bx; local.y -= by; local.z -= bz;
        const uint64_t bi = sample_index((uint32_t)c.cell[0], (uint32_t)c.cell[1], (uint32_t)c.cell[2], w.cx, w.cy);
        Level brick{w.brick_p + bi * w.words_per_brick, f, f, f};
        o.work.E++;
        DdaOut q;
        dda(brick, local, d, nullptr, 0, f, q, nullptr);
        total += q.steps; o.fine_steps += q.steps;
        hit_pos = V3{q.point.x + bx, q.point.y + by, q.point.z + bz};
        if (q.hit) {
            normal = q.steps == 0 ? c.normal : q.normal;
            hit = true;
            o.brick = (uint32_t)bi;
            for (int k = 0; k < 3; ++k) o.voxel[k] = c.cell[k] * f + q.cell[k];
            break;
        }
        start = V3{hit_pos.x / ff, hit_pos.y / ff, hit_pos.z / ff};
        if (q.oob) {
            auto same = [&](V3 p) { return hc[0] == (float)f2i(p.x) && hc[1] == (float)f2i(p.y) && hc[2] == (float)f2i(p.z); };
            if (same(start)) {
                /* push every axis that still truncates into this brick one ulp along the ray */
                const bool ex = hc[0] == (float)f2i(start.x), ey = hc[1] == (float)f2i(start.y), ez = hc[2] == (float)f2i(start.z);
                if (ex) start.x = std::nextafterf(start.x, d.x < 0 ? -kInf : kInf);
                if (ey) start.y = std::nextafterf(start.y, d.y < 0 ? -kInf : kInf);
                if (ez) start.z = std::nextafterf(start.z, d.z < 0 ? -kInf : kInf);
                if (same(start)) {
                    /* NextCell of the COARSE traversal (results.NextCell, VolumeRaytracer.cu:473) */
                    const float gx = (float)c.next[0] - start.x, gy = (float)c.next[1] - start.y, gz = (float)c.next[2] - start.z;
                    const float ax = std::fabs(gx), ay = std::fabs(gy), az = std::fabs(gz);
                    if (ax < ay && ax < az) start.x += gx;
                    else if (ay < ax && ay < az) start.y += gy;
                    else start.z += gz;
                }
            }
        }
    }
    o.steps = total;
    o.normal = normal;
    o.hit = hit;
    if (hit) {
        o.pos = hit_pos;
        if (total == 0) { o.pos = V3{start.x * ff, start.y * ff, start.z * ff}; o.normal = start_normal; }
    }
    return hit;
}

} // namespace vxo

using namespace vxo;

extern "C" int vxo_ray_aabb(const float s[3], const float d[3], const float lo[3], const float hi[3], float p[3], float n[3]) {
    V3 pp{0, 0, 0}, nn{0, 0, 0};
    bool h = ray_aabb(V3{s[0], s[1], s[2]}, V3{d[0], d[1], d[2]}, V3{lo[0], lo[1], lo[2]}, V3{hi[0], hi[1], hi[2]}, &pp, &nn);
    p[0] = pp.x; p[1] = pp.y; p[2] = pp.z; n[0] = nn.x; n[1] = nn.y; n[2] = nn.z;
    return h;
}

extern "C" void vxo_trace(const vxo_world* w, const float* origins, const float* dirs, size_t n, int norm_mode,
                          int max_steps, const vxo_hits* out, int nthreads) {
    parallel_for(n, nthreads, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            MarchOut m;
            bool h = march(*w, V3{origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]},
                           V3{dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]}, norm_mode, max_steps, m);
            if (out->hit) out->hit[i] = h;
            if (out->pos) {
                out->pos[3 * i] = h ? m.pos.x : kInf; out->pos[3 * i + 1] = h ? m.pos.y : kInf; out->pos[3 * i + 2] = h ? m.pos.z : kInf;
            }
            if (out->normal) { out->normal[3 * i] = m.normal.x; out->normal[3 * i + 1] = m.normal.y; out->normal[3 * i + 2] = m.normal.z; }
            if (out->steps) out->steps[i] = m.steps;
            if (out->voxel) for (int k = 0; k < 3; ++k) out->voxel[3 * i + k] = h ? m.voxel[k] : -1;
            if (out->brick) out->brick[i] = h ? m.brick : 0xFFFFFFFFu;
            if (out->coarse_steps) out->coarse_steps[i] = m.coarse_steps;
            if (out->fine_steps) out->fine_steps[i] = m.fine_steps;
            if (out->counters) { out->counters[3 * i] = m.work.C; out->counters[3 * i + 1] = m.work.K; out->counters[3 * i + 2] = m.work.E; }
        }
    });
}

extern "C" void vxo_trace_dense(const vxo_world* w, const float* origins, const float* dirs, size_t n, int norm_mode,
                                uint8_t* hit, int32_t* steps, int nthreads) {
    /* single-level DDARayTraversal on the dense grid: per_voxel_bounds = nullptr, bounds = nullptr
     * (BASELINE config 1 (i)) */
    Level g{w->dense.data(), (int)w->X, (int)w->Y, (int)w->Z};
    parallel_for(n, nthreads, [&](size_t lo, size_t hi) {
        for (size_t i = lo; i < hi; ++i) {
            V3 s{origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]};
            V3 r{dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]};
            V3 d = norm_mode == VXO_NORM_NONE ? r : normalize_host(r);
            if (!(s.x >= 0 && s.y >= 0 && s.z >= 0 && s.x < (float)g.nx && s.y < (float)g.ny && s.z < (float)g.nz)) {
                const float e = (float)kEpsDDA;
                V3 p, nn;
                if (ray_aabb(s, d, V3{e, e, e}, V3{(float)((double)g.nx - kEpsDDA), (float)((double)g.ny - kEpsDDA), (float)((double)g.nz - kEpsDDA)}, &p, &nn)) s = p;
            }
            DdaOut q;
            dda(g, s, d, nullptr, 0, -1, q, nullptr);
            if (hit) hit[i] = q.hit && !q.oob;
            if (steps) steps[i] = q.steps;
        }
    });
}

/* one DDA call, same contract as vxref_dda in ref_host_wrap.cu (unit tests) */
extern "C" void vxo_dda(const vxo_world* w, int level, uint32_t brick_index, const float* s, const float* d,
                        int* flags, int* cells6, float* pn6, int* steps) {
    DdaOut r;
    V3 S{s[0], s[1], s[2]}, D{d[0], d[1], d[2]};
    if (level == 0) {
        Level top{w->top_p, (int)w->cx, (int)w->cy, (int)w->cz};
        dda(top, S, D, w->bounds_p, w->factor, -1, r, nullptr);
    } else {
        Level brick{w->brick_p + (size_t)brick_index * w->words_per_brick, w->factor, w->factor, w->factor};
        dda(brick, S, D, nullptr, 0, w->factor, r, nullptr);
    }
    flags[0] = r.hit; flags[1] = r.oob; *steps = r.steps;
    for (int k = 0; k < 3; ++k) { cells6[k] = r.cell[k]; cells6[3 + k] = r.next[k]; }
    pn6[0] = r.point.x; pn6[1] = r.point.y; pn6[2] = r.point.z; pn6[3] = r.normal.x; pn6[4] = r.normal.y; pn6[5] = r.normal.z;
}
