/*
 * vxo_render.cpp — CPU ORACLE (test infrastructure, not product code): camera, pixel mapping, shading.
 *
 * Restates Renderer.cu:27-42 (GetDirections), :44-59 (getRayDirection), :72-87 (setPixelColor),
 * :90-168 (calculateColor, with the shadow Raytrace of :102 restorable), :170-177 (Tonemap),
 * :179-276 (screenDispatch: checkerboard remap, debug quadrants, miss colour, crosshair, step heat map).
 * helper_math.h semantics used: dot (:1264), length (:1307), normalize (:1325), reflect (:1427),
 * lerp (:1146), cross (:1436).
 * tanf/powf come from the host libm here and from the CUDA math library on the device: the caller
 * may pass the device's tanf value in vxo_camera::tan_half (SURVEY H1 iii); RGB parity tolerance
 * is 1/255 per channel.
 * Parity status: no executable CPU form of Renderer.cu exists (it owns __device__ globals); pinned
 * only on the GPU box against the reference kernel built by oracle/Makefile (ref_gpu).
 */
#include "vxo.h"
#include "vxo_common.h"
#include "vxo_trace.h"
#include "vxo_world.h"
#include <atomic>
#include <mutex>

using namespace vxo;

static inline float dot3(V3 a, V3 b) { float s = a.x * b.x; s = s + a.y * b.y; s = s + a.z * b.z; return s; }
static inline V3 norm3(V3 v, int mode) {
    if (mode == VXO_NORM_NONE) return v;
    return normalize_host(v);
}

extern "C" void vxo_get_directions(const float e[3], float fwd[3], float up[3], float right[3]) {
    /* cos/sin of a float argument pick the float overloads (Renderer.cu:32-37) */
    const float cxr = std::cos(e[0]), sxr = std::sin(e[0]), cyr = std::cos(e[1]), syr = std::sin(e[1]);
    const float fx = cxr * syr, fy = -sxr, fz = cxr * cyr;
    const float rx = cyr, ry = 0.0f, rz = -syr;
    const float ux = fy * rz - fz * ry, uy = fz * rx - fx * rz, uz = fx * ry - fy * rx;
    fwd[0] = fx * -1.0f; fwd[1] = fy * -1.0f; fwd[2] = fz * -1.0f;
    up[0] = ux * -1.0f; up[1] = uy * -1.0f; up[2] = uz * -1.0f;
    right[0] = rx; right[1] = ry; right[2] = rz;
}

namespace {

struct Frame {
    const vxo_world* w; const vxo_camera* cam; const vxo_env* env; const vxo_render_opts* o;
    uint32_t W, H; uint8_t* bgra; int32_t* steps_plane;
    float tan_half;
};
/* per-thread accumulators of vxo_render_sample: primary rays, primary hits, shadow rays, shadow hits,
 * steps, C, K, E (same order as vx_render_opts::d_stats) */
struct Tally { uint64_t v[8] = {0, 0, 0, 0, 0, 0, 0, 0}; };

void put(const Frame& f, int x, int y, V3 c) {
    if (!f.bgra || (uint32_t)x >= f.W || (uint32_t)y >= f.H) return;
    uint8_t* p = f.bgra + ((size_t)y * f.W + x) * 4;
    c.x = fmin_cu(fmax_cu(c.x, 0.0f), 1.0f);
    c.y = fmin_cu(fmax_cu(c.y, 0.0f), 1.0f);
    c.z = fmin_cu(fmax_cu(c.z, 0.0f), 1.0f);
    p[2] = (uint8_t)f2i(c.x * 255.0f);
    p[1] = (uint8_t)f2i(c.y * 255.0f);
    p[0] = (uint8_t)f2i(c.z * 255.0f);
    p[3] = 255;
}

V3 primary_dir(const Frame& f, int x, int y) {
    const float u0 = (float)x / (float)f.W, v0 = (float)y / (float)f.H;
    const float aspect = (float)f.W / (float)f.H;
    const float u = u0 * 2.0f - 1.0f, v = v0 * 2.0f - 1.0f;
    const float sx = f.tan_half * aspect, sy = f.tan_half;
    const V3 F{f.cam->fwd[0], f.cam->fwd[1], f.cam->fwd[2]}, U{f.cam->up[0], f.cam->up[1], f.cam->up[2]},
        R{f.cam->right[0], f.cam->right[1], f.cam->right[2]};
    const float a = u * sx, b = v * sy;
    V3 d;
    d.x = F.x + a * R.x + b * U.x;
    d.y = F.y + a * R.y + b * U.y;
    d.z = F.z + a * R.z + b * U.z;
    return norm3(d, f.o->norm_mode);
}

/* getRayDirectionOrtho, Renderer.cu:61-70: direction = fwd as given (Raytrace normalises it, once), the origin moves
 * across the window: origin += right * (u*2-1) * size.x * ratio; origin += up * (v*2-1) * size.y — float3 * float
 * products left to right, each a rounded float (helper_math operator*, operator+=) */
bool ortho(const Frame& f) { return f.cam->ortho_w > 0.0f && f.cam->ortho_h > 0.0f; }
V3 ortho_origin(const Frame& f, int x, int y) {
    const float u0 = (float)x / (float)f.W, v0 = (float)y / (float)f.H;
    const float ratio = (float)f.W / (float)f.H;
    const float a = u0 * 2.0f - 1.0f, b = v0 * 2.0f - 1.0f;
    const V3 U{f.cam->up[0], f.cam->up[1], f.cam->up[2]}, R{f.cam->right[0], f.cam->right[1], f.cam->right[2]};
    V3 o{f.cam->origin[0], f.cam->origin[1], f.cam->origin[2]};
    V3 t{R.x * a, R.y * a, R.z * a};
    t = V3{t.x * f.cam->ortho_w, t.y * f.cam->ortho_w, t.z * f.cam->ortho_w};
    t = V3{t.x * ratio, t.y * ratio, t.z * ratio};
    o = V3{o.x + t.x, o.y + t.y, o.z + t.z};
    t = V3{U.x * b, U.y * b, U.z * b};
    t = V3{t.x * f.cam->ortho_h, t.y * f.cam->ortho_h, t.z * f.cam->ortho_h};
    o = V3{o.x + t.x, o.y + t.y, o.z + t.z};
    return o;
}

V3 shade(const Frame& f, V3 cam, V3 n, V3 pos, int& extra_steps, Tally* tl) {
    const V3 L{f.env->light_dir[0], f.env->light_dir[1], f.env->light_dir[2]};
    const V3 LC{f.env->light_color[0], f.env->light_color[1], f.env->light_color[2]};
    const V3 AC{f.env->ambient[0], f.env->ambient[1], f.env->ambient[2]};
    extra_steps = 0;
    const V3 sray = norm3(L, f.o->norm_mode);
    const V3 spos{pos.x + sray.x * 0.01f, pos.y + sray.y * 0.01f, pos.z + sray.z * 0.01f};
    bool shadowed = false;
    if (f.o->shadows) {
        MarchOut m;
        shadowed = march(*f.w, spos, sray, f.o->norm_mode, kMaxSteps, m);
        extra_steps += m.steps;
        if (tl) { tl->v[2] += 1; tl->v[3] += shadowed; tl->v[5] += m.work.C; tl->v[6] += m.work.K; tl->v[7] += m.work.E; }
    }
    const float ldot = fmax_cu(dot3(n, L), 0.0f) * (shadowed ? 0.0f : 1.0f);
    const float up_dot = dot3(n, V3{0, 1, 0});
    const float t = (float)((double)up_dot * 0.5 + 0.5); /* double expression narrowed to lerp's float t */
    const float amb = 0.25f + t * (1.0f - 0.25f);
    V3 c{ldot * LC.x + AC.x * amb, ldot * LC.y + AC.y * amb, ldot * LC.z + AC.z * amb};
    if (!shadowed) {
        const V3 view = norm3(V3{pos.x - cam.x, pos.y - cam.y, pos.z - cam.z}, f.o->norm_mode);
        const float nl = dot3(n, L);
        const V3 refl{L.x - 2.0f * n.x * nl, L.y - 2.0f * n.y * nl, L.z - 2.0f * n.z * nl};
        const float spec = std::pow(fmax_cu(dot3(view, refl), 0.0f), 32.0f);
        c.x += spec * LC.x; c.y += spec * LC.y; c.z += spec * LC.z;
    }
    /* AO block: samples = 0 -> occlusion = 1.0f, colour unchanged (Renderer.cu:121-165) */
    return c;
}

void pixel(const Frame& f, int tx, int ty, Tally* tl = nullptr) {
    int x = tx, y = ty;
    if (f.o->checkerboard) {
        y *= 2;
        if ((x % 2) == 0) y += 1;
        if (f.o->frame_number % 2 == 0) y += 1;
    }
    if ((uint32_t)x >= f.W || (uint32_t)y >= f.H) return;
    const int W = (int)f.W, H = (int)f.H;
    const bool orth = ortho(f);
    const V3 origin = orth ? ortho_origin(f, x, y) : V3{f.cam->origin[0], f.cam->origin[1], f.cam->origin[2]};
    const V3 rd = orth ? V3{f.cam->fwd[0], f.cam->fwd[1], f.cam->fwd[2]} : primary_dir(f, x, y);
    MarchOut m;
    const bool hit = march(*f.w, origin, rd, f.o->norm_mode, kMaxSteps, m);
    int steps = m.steps;
    if (tl) { tl->v[0] += 1; tl->v[1] += hit; tl->v[5] += m.work.C; tl->v[6] += m.work.K; tl->v[7] += m.work.E; }
    const V3 n{-m.normal.x, -m.normal.y, -m.normal.z};
    if (hit) {
        if (f.o->debug_view) {
            const V3 dlt{m.pos.x - origin.x, m.pos.y - origin.y, m.pos.z - origin.z};
            const float dist = std::sqrt(dot3(dlt, dlt));
            const float md = (float)(1.0f + kEpsDDA);
            const V3 hp{std::fmod(m.pos.x / 128.0f, md), std::fmod(m.pos.y / 128.0f, md), std::fmod(m.pos.z / 128.0f, md)};
            if (x < (W >> 1) && y < (H >> 1)) put(f, x, y, n);
            else if (x >= (W >> 1) && y < (H >> 1)) put(f, x, y, hp);
            else if (x < (W >> 1)) { /* bottom left: nothing */ }
            else put(f, x, y, V3{dist * 0.01f, 0, 0});
        } else {
            int extra = 0;
            V3 c = shade(f, origin, n, m.pos, extra, tl);
            c = V3{c.x / (c.x + 1.0f), c.y / (c.y + 1.0f), c.z / (c.z + 1.0f)};
            c.x = fmin_cu(fmax_cu(c.x, 0.0f), 1.0f); c.y = fmin_cu(fmax_cu(c.y, 0.0f), 1.0f); c.z = fmin_cu(fmax_cu(c.z, 0.0f), 1.0f);
            steps += extra;
            put(f, x, y, c);
        }
    } else {
        put(f, x, y, rd);
    }
    if (f.o->crosshair && (uint32_t)tx == (f.W >> 1) && (uint32_t)ty == (f.H >> 1)) put(f, x, y, V3{10, 10, 10});
    if (f.o->debug_view && x < (W >> 1) && y > (H >> 1)) put(f, x, y, V3{(float)steps / 256.0f, 0, 0});
    if (f.steps_plane) f.steps_plane[(size_t)y * f.W + x] = steps;
    if (tl) tl->v[4] += (uint64_t)steps;
}

} // namespace

extern "C" void vxo_render(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                           uint32_t W, uint32_t H, uint8_t* bgra, int32_t* steps_plane, int nthreads) {
    Frame f{w, cam, env, o, W, H, bgra, steps_plane, 0.0f};
    if (cam->tan_half > 0.0f) f.tan_half = cam->tan_half;
    else {
        const float fov = (float)((double)cam->fov_deg * 3.1415 / 180.0);
        f.tan_half = std::tan(fov / 2.0f); /* float overload == tanf */
    }
    const uint32_t rows = o->checkerboard ? (H >> 1) : H; /* grid.y of RenderScreen, Renderer.cu:311-316 */
    const uint32_t cols = (W + 31) / 32 * 32;
    parallel_for(rows, nthreads, [&](size_t lo, size_t hi) {
        for (size_t ty = lo; ty < hi; ++ty)
            for (uint32_t tx = 0; tx < cols; ++tx) pixel(f, (int)tx, (int)ty);
    });
}

/* vxo_render_sample with the two planes a parity check needs: pixels of the sampled thread rows are written to bgra (W*H*4
 * bytes, others untouched) and their step counts (primary + sun-ray DDA steps) to steps_plane (W*H int32). */
extern "C" void vxo_render_sample_planes(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                                         uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                                         uint8_t* bgra, int32_t* steps_plane, int nthreads) {
    Frame f{w, cam, env, o, W, H, bgra, steps_plane, 0.0f};
    if (cam->tan_half > 0.0f) f.tan_half = cam->tan_half;
    else {
        const float fov = (float)((double)cam->fov_deg * 3.1415 / 180.0);
        f.tan_half = std::tan(fov / 2.0f);
    }
    const uint32_t rows = o->checkerboard ? (H >> 1) : H;
    if (ty_end > rows) ty_end = rows;
    if (ty_stride == 0) ty_stride = 1;
    const uint32_t cols = (W + 31) / 32 * 32;
    const size_t nrows = ty_end > ty_begin ? (ty_end - ty_begin + ty_stride - 1) / ty_stride : 0;
    parallel_for(nrows, nthreads, [&](size_t lo, size_t hi) {
        for (size_t r = lo; r < hi; ++r)
            for (uint32_t tx = 0; tx < cols; ++tx) pixel(f, (int)tx, (int)(ty_begin + r * ty_stride));
    });
}

/* The rays screenDispatch hands to Raytrace for the sampled thread rows, as a flat list for a batch tracer (bench.py's
 * reference arm feeds them to the reference's own Raytrace compiled for the host): every pixel's primary ray
 * (Renderer.cu:199-211) and, for every primary hit, calculateColor's shadow ray (:98-102).  origins/dirs hold up to
 * `capacity` rays of 3 floats; returns the number of rays (which may exceed capacity: nothing is written past it). */
extern "C" size_t vxo_frame_rays(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                                 uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                                 float* origins, float* dirs, size_t capacity) {
    Frame f{w, cam, env, o, W, H, nullptr, nullptr, 0.0f};
    if (cam->tan_half > 0.0f) f.tan_half = cam->tan_half;
    else {
        const float fov = (float)((double)cam->fov_deg * 3.1415 / 180.0);
        f.tan_half = std::tan(fov / 2.0f);
    }
    const uint32_t rows = o->checkerboard ? (H >> 1) : H;
    if (ty_end > rows) ty_end = rows;
    if (ty_stride == 0) ty_stride = 1;
    const uint32_t cols = (W + 31) / 32 * 32;
    const V3 L{env->light_dir[0], env->light_dir[1], env->light_dir[2]};
    const V3 sray = norm3(L, o->norm_mode);
    const size_t nrows = ty_end > ty_begin ? (ty_end - ty_begin + ty_stride - 1) / ty_stride : 0;
    std::vector<std::vector<float>> per_row(nrows); /* 6 floats per ray, rows marched in parallel, order kept */
    parallel_for(nrows, 0, [&](size_t lo, size_t hi) {
        for (size_t r = lo; r < hi; ++r) {
            std::vector<float>& out = per_row[r];
            const uint32_t ty = ty_begin + (uint32_t)r * ty_stride;
            for (uint32_t tx = 0; tx < cols; ++tx) {
                int x = (int)tx, y = (int)ty;
                if (o->checkerboard) { y *= 2; if ((x % 2) == 0) y += 1; if (o->frame_number % 2 == 0) y += 1; }
                if ((uint32_t)x >= W || (uint32_t)y >= H) continue;
                const bool orth = ortho(f);
                const V3 origin = orth ? ortho_origin(f, x, y) : V3{cam->origin[0], cam->origin[1], cam->origin[2]};
                const V3 rd = orth ? V3{cam->fwd[0], cam->fwd[1], cam->fwd[2]} : primary_dir(f, x, y);
                out.insert(out.end(), {origin.x, origin.y, origin.z, rd.x, rd.y, rd.z});
                if (o->debug_view || !o->shadows) continue;
                MarchOut m;
                if (march(*w, origin, rd, o->norm_mode, kMaxSteps, m))
                    out.insert(out.end(), {m.pos.x + sray.x * 0.01f, m.pos.y + sray.y * 0.01f, m.pos.z + sray.z * 0.01f, sray.x, sray.y, sray.z});
            }
        }
    });
    size_t n = 0;
    for (const auto& row : per_row)
        for (size_t k = 0; k + 5 < row.size(); k += 6, ++n)
            if (n < capacity) { std::memcpy(origins + 3 * n, &row[k], 12); std::memcpy(dirs + 3 * n, &row[k + 3], 12); }
    return n;
}

/* A bounded sample of a frame for the timed CPU baseline: thread rows ty_begin, ty_begin+ty_stride, ...
 * < ty_end (clamped to the frame's grid).  bgra may be NULL (nothing is stored).  stats_out[8] receives
 * primary rays, primary hits, shadow rays, shadow hits, steps, C, K, E of the sample. */
extern "C" void vxo_render_sample(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                                  uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                                  uint8_t* bgra, uint64_t* stats_out, int nthreads) {
    Frame f{w, cam, env, o, W, H, bgra, nullptr, 0.0f};
    if (cam->tan_half > 0.0f) f.tan_half = cam->tan_half;
    else {
        const float fov = (float)((double)cam->fov_deg * 3.1415 / 180.0);
        f.tan_half = std::tan(fov / 2.0f);
    }
    const uint32_t rows = o->checkerboard ? (H >> 1) : H;
    if (ty_end > rows) ty_end = rows;
    if (ty_stride == 0) ty_stride = 1;
    const uint32_t cols = (W + 31) / 32 * 32;
    const size_t nrows = ty_end > ty_begin ? (ty_end - ty_begin + ty_stride - 1) / ty_stride : 0;
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    std::vector<Tally> tallies((size_t)nthreads * 16 + 1);
    std::atomic<size_t> slot{0};
    parallel_for(nrows, nthreads, [&](size_t lo, size_t hi) {
        Tally local;
        for (size_t r = lo; r < hi; ++r)
            for (uint32_t tx = 0; tx < cols; ++tx) pixel(f, (int)tx, (int)(ty_begin + r * ty_stride), &local);
        Tally& dst = tallies[slot.fetch_add(1) % tallies.size()];
        static std::mutex mu;
        std::lock_guard<std::mutex> g(mu);
        for (int k = 0; k < 8; ++k) dst.v[k] += local.v[k];
    });
    for (int k = 0; k < 8; ++k) { stats_out[k] = 0; for (auto& t : tallies) stats_out[k] += t.v[k]; }
}
