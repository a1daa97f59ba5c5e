// vx_trace.cu — batch ray kernel, screen render kernel and their C ABI.
//
// Replaces the reference's `dispatch` kernel + VoxelRaytracer3D::Raytrace (VolumeRaytracer.cu:95-117,
// :574-618) and `screenDispatch` + Graphics::RenderScreen (Renderer.cu:179-328).  The march itself is
// vx_march.cuh (per-lane state machine, persistent warps); this file holds the two ray sources/sinks
// ("policies") that feed it, the launches and the C ABI.
#include "vx_host.h"
#include "vx_march.cuh"
#include "vx_noise.cuh"
#include <stdlib.h>
#include <string.h>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

#ifndef VX_TILE_SHAPE
#define VX_TILE_SHAPE 0 // pixels a warp of primary_rays takes at a time: 0 = 8x4, 1 = 16x2, 2 = 32x1
#endif
#ifndef VX_TRACE_MIN_BLOCKS
#define VX_TRACE_MIN_BLOCKS 10 // the batch kernel: 10 since round 2 (bounce rays 2.11 -> 2.05 ms; 8: 2.11, 9: 2.10, 11: 2.07, 12: 2.08)
#endif
#ifndef VX_SHADOW_MIN_BLOCKS
#define VX_SHADOW_MIN_BLOCKS 10 // sun rays: 10 since the chain trims of round 2 (0.840 -> 0.825 ms; 11, 12: the same with spills; round 1: 9)
#endif
#ifndef VX_MIN_BLOCKS
#define VX_MIN_BLOCKS 9 // resident 128-thread blocks per SM the render kernels are compiled for (56-register budget; measured
                        // per 4K frame: 6 -> 2.96 ms, 7 -> 2.88, 8 -> 2.85, 9 -> 2.83, 10 (shadow_rays) -> 2.89)
#endif

namespace vx {

struct HitPlanes {
    uint8_t* hit; float* pos; float* normal; int32_t* steps; int32_t* voxel; uint32_t* brick;
    int32_t* coarse_steps; int32_t* fine_steps; uint32_t* counters;
};

// ---- batch rays ---------------------------------------------------------------------------------
template <int NORM, bool COUNT>
struct BatchPolicy {
    const float* __restrict__ origins;
    const float* __restrict__ dirs;
    unsigned long long n;
    int cap;
    HitPlanes out;
    unsigned long long id; // per lane: the ray being marched
    static constexpr int kRefill = VX_REFILL_MIN;
    static constexpr bool kDrain = false, kSun = false;
    static constexpr int kFastQuad = VX_BATCH_QUAD, kLeaveNum = VX_BATCH_LEAVE_NUM, kLeaveDen = VX_BATCH_LEAVE_DEN;

    __device__ __forceinline__ unsigned long long count() const { return n; }
    __device__ __forceinline__ int max_steps() const { return cap; }
    __device__ __forceinline__ void prepare(Lane&) {}
    __device__ __forceinline__ void flush() {}
    __device__ __forceinline__ void init(const WorldView& w, unsigned long long i, Lane& L) {
        id = i;
        const float ff = (float)w.factor;
        const V3 o = mk(origins[3 * i], origins[3 * i + 1], origins[3 * i + 2]);
        const V3 d = normalize3<NORM>(mk(dirs[3 * i], dirs[3 * i + 1], dirs[3 * i + 2]));
        lane_setup(w, L, to_brick_units(w, o), d, 0u);
    }
    __device__ __forceinline__ void finish(const WorldView& w, Lane& L, V3 pos, uint32_t ncode) {
        const unsigned long long i = id;
        const bool h = L.flags & F_HIT;
        if (out.hit) out.hit[i] = h;
        if (out.pos) {
            out.pos[3 * i] = h ? pos.x : CUDART_INF_F; out.pos[3 * i + 1] = h ? pos.y : CUDART_INF_F; out.pos[3 * i + 2] = h ? pos.z : CUDART_INF_F;
        }
        if (out.normal) {
            const V3 nn = normal_from_code(h ? ncode : 0u);
            out.normal[3 * i] = nn.x; out.normal[3 * i + 1] = nn.y; out.normal[3 * i + 2] = nn.z;
        }
        if (out.steps) out.steps[i] = L.total;
        if (out.voxel || out.brick) {
            int bx = 0, by = 0, bz = 0;
            if (h) cell_of(w, L.pc, bx, by, bz);
            if (out.voxel) {
                out.voxel[3 * i] = h ? bx * w.factor + (int)(L.fcell & 1023u) : -1;
                out.voxel[3 * i + 1] = h ? by * w.factor + (int)((L.fcell >> 10) & 1023u) : -1;
                out.voxel[3 * i + 2] = h ? bz * w.factor + (int)((L.fcell >> 20) & 1023u) : -1;
            }
            if (out.brick) out.brick[i] = h ? tiled_index(bx, by, bz, w.tiles_x, w.tiles_xy) : 0xFFFFFFFFu;
        }
        if (out.coarse_steps) out.coarse_steps[i] = L.csteps;
        if (out.fine_steps) out.fine_steps[i] = L.fsteps;
        if (COUNT && out.counters) { out.counters[3 * i] = L.work.C; out.counters[3 * i + 1] = L.work.K; out.counters[3 * i + 2] = L.work.E; }
        L.flags = ST_IDLE;
    }
};

// F8 (every march kernel): the world's bricks are 8^3, known at compile time.  The kernels for any other factor carry the generic
// brick march, nine IEEE divisions by the factor and the loads of the three factor-dependent constants; the march kernels are
// instruction-fetch sensitive (58 KB of code, `no_instruction` stalls), so the common case gets kernels without that code.
__device__ __forceinline__ WorldView factor8_view(WorldView w) {
    w.factor = 8; w.inv_factor = 0.125f; w.words_per_brick = 16u;
    return w;
}

template <int NORM, bool COUNT, bool F8>
__global__ void __launch_bounds__(128, VX_TRACE_MIN_BLOCKS) trace_rays(WorldView w_in, const float* __restrict__ origins, const float* __restrict__ dirs,
                                                  unsigned long long n, int max_steps, HitPlanes out, unsigned long long* next) {
    const WorldView w = F8 ? factor8_view(w_in) : w_in;
    BatchPolicy<NORM, COUNT> pol;
    pol.origins = origins; pol.dirs = dirs; pol.n = n; pol.cap = max_steps; pol.out = out; pol.id = 0;
    run_lanes<true, COUNT>(w, pol, next);
}

// ---- screen ----------------------------------------------------------------------------------
struct FrameParams {
    V3 origin, fwd, up, right;
    V3 start0;                 // origin / factor (IEEE division, done once on the host)
    float ortho_w, ortho_h;    // > 0: orthographic window (getRayDirectionOrtho, Renderer.cu:61-70), else perspective
    float factor_f;
    V3 light, light_color, ambient;
    float tan_half;
    uint32_t W, H;
    uint32_t frame_number;
    int checkerboard, debug_view, shadows, crosshair;
    uint32_t tx0, ty0, tw, th; // thread-space rectangle
    uint32_t band_count, band_index; // multi-GPU: 8-row bands dealt round-robin, this launch renders bands == index (mod count)
    uint32_t mcols;            // 16x8 macro tiles per band
    unsigned long long ids;    // ray ids of this launch = mcols * my_bands * 128
    uint32_t* __restrict__ tile_trace; // diagnostics (VX_TILE_TRACE=1), else NULL: per 8x4 tile of the launch, the SM cycles it kept its
                                       // warp and (second half of the array) the low word of %globaltimer when the warp drew it
};

__device__ __forceinline__ void put_pixel(uint8_t* __restrict__ bgra, uint32_t W, uint32_t H, int x, int y, V3 c) {
    // setPixelColor, Renderer.cu:72-87: clamp, float->uint8 truncation, B,G,R,A byte order
    if ((uint32_t)x >= W || (uint32_t)y >= H) return;
    c.x = fminf(fmaxf(c.x, 0.0f), 1.0f); c.y = fminf(fmaxf(c.y, 0.0f), 1.0f); c.z = fminf(fmaxf(c.z, 0.0f), 1.0f);
    const uint32_t r = (uint32_t)(int)fmul(c.x, 255.0f), g = (uint32_t)(int)fmul(c.y, 255.0f), b = (uint32_t)(int)fmul(c.z, 255.0f);
    reinterpret_cast<uint32_t*>(bgra)[(size_t)y * W + x] = b | (g << 8) | (r << 16) | 0xFF000000u;
}

// The origin screenDispatch traces pixel (x, y) from, and shades with: the camera, or in the ORTHO build (Renderer.cu:202-204)
// getRayDirectionOrtho's origin += right * (u*2-1) * size.x * ratio; origin += up * (v*2-1) * size.y (:61-70; float3 * float
// products left to right)
__device__ __forceinline__ V3 pixel_origin(const FrameParams& f, bool ortho, int x, int y) {
    if (!ortho) return f.origin;
    const float a = fsub(fmul(fdiv((float)x, (float)(int)f.W), 2.0f), 1.0f), b = fsub(fmul(fdiv((float)y, (float)(int)f.H), 2.0f), 1.0f);
    const float ratio = fdiv((float)f.W, (float)f.H);
    V3 t = mk(fmul(fmul(fmul(f.right.x, a), f.ortho_w), ratio), fmul(fmul(fmul(f.right.y, a), f.ortho_w), ratio), fmul(fmul(fmul(f.right.z, a), f.ortho_w), ratio));
    V3 o = mk(fadd(f.origin.x, t.x), fadd(f.origin.y, t.y), fadd(f.origin.z, t.z));
    t = mk(fmul(fmul(f.up.x, b), f.ortho_h), fmul(fmul(f.up.y, b), f.ortho_h), fmul(fmul(f.up.z, b), f.ortho_h));
    return mk(fadd(o.x, t.x), fadd(o.y, t.y), fadd(o.z, t.z));
}

// calculateColor, Renderer.cu:90-168 (AO block: samples = 0, no-op) with the shadow result supplied, then Tonemap (:170-177)
template <int NORM>
__device__ __forceinline__ V3 shade(const FrameParams& f, V3 cam, V3 n, V3 pos, bool shadowed) {
    const float nl = dot3(n, f.light);
    const float ldot = fmul(fmaxf(nl, 0.0f), shadowed ? 0.0f : 1.0f);
    const float up_dot = dot3(n, mk(0.0f, 1.0f, 0.0f));
    const float t = (float)__dadd_rn(__dmul_rn((double)up_dot, 0.5), 0.5); // double expression, Renderer.cu:106
    const float amb = fadd(0.25f, fmul(t, 0.75f));                          // helper_math lerp(0.25, 1.0, t)
    V3 c = mk(fadd(fmul(ldot, f.light_color.x), fmul(f.ambient.x, amb)), fadd(fmul(ldot, f.light_color.y), fmul(f.ambient.y, amb)),
              fadd(fmul(ldot, f.light_color.z), fmul(f.ambient.z, amb)));
    if (!shadowed) {
        const V3 view = normalize3<NORM>(mk(fsub(pos.x, cam.x), fsub(pos.y, cam.y), fsub(pos.z, cam.z)));
        const V3 refl = mk(fsub(f.light.x, fmul(fmul(2.0f, n.x), nl)), fsub(f.light.y, fmul(fmul(2.0f, n.y), nl)),
                           fsub(f.light.z, fmul(fmul(2.0f, n.z), nl)));
        const float spec = powf(fmaxf(dot3(view, refl), 0.0f), 32.0f);
        c = mk(fadd(c.x, fmul(spec, f.light_color.x)), fadd(c.y, fmul(spec, f.light_color.y)), fadd(c.z, fmul(spec, f.light_color.z)));
    }
    return mk(fdiv(c.x, fadd(c.x, 1.0f)), fdiv(c.y, fadd(c.y, 1.0f)), fdiv(c.z, fadd(c.z, 1.0f)));
}

// A shadowed frame is four launches:
//   primary_rays : persistent warps, one 8x4 pixel tile per refill (coherent rays).  A miss writes its pixel; a hit
//                  is either shaded at once (shadows off / debug view) or appended to the frame's shadow queue.
//   shadow_setup : dense, one thread per queued hit: the sun ray's set-up, its first brick and the traversal start after it.
//   shadow_rays  : persistent warps pull the lane records (neighbouring pixels sit next to each other in the queue) and
//                  refill finished lanes in groups, because about half of the sun rays end after a few steps while
//                  the rest cross the whole grid.
//   shade_pixels : dense: entry + result -> pixel (or, FUSE, inside shadow_rays for a frame in another GPU's memory).
// Ray id -> pixel for the primary launch: ids are dealt in 8x4 tiles (32 consecutive ids), four tiles forming a
// 16x8 macro tile, macro tiles row-major inside this launch's bands.  Pixel (thread) coordinates and the written
// image are the reference's.
//
// STATS: per-frame counters (8 x uint64: primary rays, primary hits, shadow rays, shadow hits, steps, C, K, E —
// SURVEY.md §8d), added with one atomic per group of lanes finishing together.
struct ShadowQueue {
    float4* __restrict__ rec;          // primary hit: position xyz (voxel units), w = pixel bits (x | y << 15 | crosshair << 30)
    uint32_t* __restrict__ aux;        // planes of `ids` words: 0 primary steps | normal code << 16, 1 sun-ray steps | hit << 31,
                                       // with STATS also 2-4 primary C,K,E and 5-7 sun-ray C,K,E
    uint4* __restrict__ lanes;         // 3 x uint4 per entry: the sun ray's march state after set-up and its first
                                       // traversal start (s.xyz tx | ty tz idx total | pc flags 0 0), written by shadow_setup
    unsigned long long* __restrict__ count;
};

__device__ __forceinline__ void add_stats(unsigned long long* __restrict__ stats, uint32_t rays, uint32_t hits, uint32_t srays,
                                          uint32_t shits, uint32_t steps, Work wk) {
    const uint32_t m = __activemask();
    const uint32_t leader = (uint32_t)__ffs((int)m) - 1u, lane = threadIdx.x & 31u;
    const uint32_t v[8] = {rays, hits, srays, shits, steps, wk.C, wk.K, wk.E};
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t sum = __reduce_add_sync(m, v[k]);
        if (lane == leader && sum) atomicAdd(stats + k, (unsigned long long)sum);
    }
}

// Everything screenDispatch does once the rays of a pixel are known (Renderer.cu:203-275).
// PLAIN: the caller only ever has misses to write and the debug view is off (primary_rays of a shadowed frame): no shading code.
template <int NORM, bool PLAIN = false>
__device__ __forceinline__ void emit_pixel(const FrameParams& f, bool ortho, uint8_t* __restrict__ bgra, int32_t* __restrict__ steps_plane, uint32_t pix,
                                           bool hit, V3 pos, uint32_t ncode, bool shadowed, int steps, V3 miss_dir) {
    const int x = (int)(pix & 32767u), y = (int)((pix >> 15) & 32767u);
    const int W = (int)f.W, H = (int)f.H;
    if (!PLAIN && hit) {
        const V3 nn = normal_from_code(ncode);
        const V3 n = mk(-nn.x, -nn.y, -nn.z);
        const V3 cam = pixel_origin(f, ortho, x, y);
        if (f.debug_view) { // Renderer.cu:215-243
            const V3 dl = mk(fsub(pos.x, cam.x), fsub(pos.y, cam.y), fsub(pos.z, cam.z));
            const float dist = __fsqrt_rn(dot3(dl, dl));
            const float md = (float)(1.0 + kEpsDDA);
            const V3 hp = mk(fmodf(fdiv(pos.x, 128.0f), md), fmodf(fdiv(pos.y, 128.0f), md), fmodf(fdiv(pos.z, 128.0f), md));
            if (x < (W >> 1) && y < (H >> 1)) put_pixel(bgra, f.W, f.H, x, y, n);
            else if (x >= (W >> 1) && y < (H >> 1)) put_pixel(bgra, f.W, f.H, x, y, hp);
            else if (x < (W >> 1)) {}
            else put_pixel(bgra, f.W, f.H, x, y, mk(fmul(dist, 0.01f), 0.0f, 0.0f));
        } else {
            put_pixel(bgra, f.W, f.H, x, y, shade<NORM>(f, cam, n, pos, shadowed));
        }
    } else {
        put_pixel(bgra, f.W, f.H, x, y, miss_dir); // miss: colour = ray direction, Renderer.cu:254-258
    }
    if (pix & (1u << 30)) put_pixel(bgra, f.W, f.H, x, y, mk(10.0f, 10.0f, 10.0f)); // crosshair, :260-268
    if (!PLAIN && f.debug_view && x < (W >> 1) && y > (H >> 1)) put_pixel(bgra, f.W, f.H, x, y, mk(fdiv((float)steps, 256.0f), 0.0f, 0.0f));
    if (steps_plane) steps_plane[(size_t)y * f.W + x] = steps;
}

// QUEUED: the launch is the first of a shadowed frame (shadows on, debug view off), known at compile time: every hit goes to
// the frame's queue and this kernel shades nothing.
template <int NORM, bool STATS, bool ORTHO, bool QUEUED>
struct PrimaryPolicy {
    static constexpr bool kDrain = false, kSun = false;
    static constexpr int kFastQuad = VX_PRIMARY_QUAD, kLeaveNum = VX_PRIMARY_LEAVE_NUM, kLeaveDen = VX_PRIMARY_LEAVE_DEN;
    static constexpr int kRefill = 32; // a warp takes a whole tile at a time: neighbouring rays stay in step (2.2 ms instead of
                                       // 1.56 ms per 4K frame with 16..28), and flush() relies on it
#ifndef VX_SHADOW_REFILL
#define VX_SHADOW_REFILL VX_REFILL_MIN
#endif
    const FrameParams& f;
    uint8_t* __restrict__ bgra;
    int32_t* __restrict__ steps_plane;
    unsigned long long* __restrict__ stats;
    ShadowQueue q;
    uint32_t pix; // per lane: x | y << 15 | crosshair thread << 30 | hit staged << 31
    float4* stage_rec; uint32_t* stage_aux; // shared memory, one slot per thread (x4 words with STATS)
    uint32_t* tile_slot;                    // shared memory, two words per warp (tile trace): the tile this warp runs, the clock at its draw

    __device__ __forceinline__ PrimaryPolicy(const FrameParams& fp, uint8_t* b, int32_t* sp, unsigned long long* st, ShadowQueue sq)
        : f(fp), bgra(b), steps_plane(sp), stats(st), q(sq), pix(0) {}
    __device__ __forceinline__ unsigned long long count() const { return f.ids; }
    __device__ __forceinline__ int max_steps() const { return kMaxSteps; }
    __device__ __forceinline__ void prepare(Lane&) {}
    // The hits of a tile enter the frame's queue together and in pixel order, whenever the warp has finished the tile:
    // the sun rays that shadow_rays later draws 32 at a time are then neighbours on the screen (coherent), which they are
    // not if every lane appends the moment it finishes.
    __device__ __forceinline__ void flush() {
        const uint32_t lane = threadIdx.x & 31u;
        if (f.tile_trace && lane == 0u && tile_slot[0] != 0xFFFFFFFFu) {
            f.tile_trace[tile_slot[0]] = (uint32_t)clock64() - tile_slot[1];
            tile_slot[0] = 0xFFFFFFFFu;
        }
        const bool staged = pix >> 31;
        const uint32_t m = __ballot_sync(kFull, staged);
        if (!m) return;
        unsigned long long base = 0;
        if (lane == 0) base = atomicAdd(q.count, (unsigned long long)__popc(m));
        base = __shfl_sync(kFull, base, 0);
        if (staged) {
            const unsigned long long slot = base + (unsigned long long)__popc(m & ((1u << lane) - 1u));
            const uint32_t tid = threadIdx.x;
            q.rec[slot] = stage_rec[tid];
            q.aux[slot] = stage_aux[tid];
            if (STATS) { q.aux[slot + 2 * f.ids] = stage_aux[tid + 128]; q.aux[slot + 3 * f.ids] = stage_aux[tid + 256]; q.aux[slot + 4 * f.ids] = stage_aux[tid + 384]; }
            pix &= 0x7FFFFFFFu;
        }
    }

    // getRayDirection, Renderer.cu:44-59
    __device__ __forceinline__ V3 primary_dir(int x, int y) const {
        const int W = (int)f.W, H = (int)f.H;
        const float u = fsub(fmul(fdiv((float)x, (float)W), 2.0f), 1.0f), v = fsub(fmul(fdiv((float)y, (float)H), 2.0f), 1.0f);
        const float aspect = fdiv((float)W, (float)H);
        const float a = fmul(u, fmul(f.tan_half, aspect)), b = fmul(v, f.tan_half);
        const V3 rd = mk(fadd(fadd(f.fwd.x, fmul(a, f.right.x)), fmul(b, f.up.x)), fadd(fadd(f.fwd.y, fmul(a, f.right.y)), fmul(b, f.up.y)),
                         fadd(fadd(f.fwd.z, fmul(a, f.right.z)), fmul(b, f.up.z)));
        return normalize3<NORM>(rd);
    }

    __device__ __forceinline__ void init(const WorldView& w, unsigned long long id64, Lane& L) {
        const uint32_t id = (uint32_t)id64;
        if (f.tile_trace && (id & 31u) == 0u) {
            unsigned long long ns;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(ns));
            tile_slot[0] = id >> 5; tile_slot[1] = (uint32_t)clock64(); f.tile_trace[(f.ids >> 5) + (id >> 5)] = (uint32_t)ns;
        }
        const uint32_t j = id & 31u, k = (id >> 5) & 3u, macro = id >> 7;
        const uint32_t lband = macro / f.mcols, mcol = macro - lband * f.mcols;
#if VX_TILE_SHAPE == 0   // four 8x4 tiles per 16x8 macro tile
        const uint32_t tx = f.tx0 + mcol * 16u + (k & 1u) * 8u + (j & 7u);
        const uint32_t ty = f.ty0 + (lband * f.band_count + f.band_index) * 8u + (k >> 1) * 4u + (j >> 3);
#elif VX_TILE_SHAPE == 1 // four 16x2 tiles per 16x8 macro tile
        const uint32_t tx = f.tx0 + mcol * 16u + (j & 15u);
        const uint32_t ty = f.ty0 + (lband * f.band_count + f.band_index) * 8u + k * 2u + (j >> 4);
#elif VX_TILE_SHAPE == 3 // four 4x8 tiles per 16x8 macro tile
        const uint32_t tx = f.tx0 + mcol * 16u + k * 4u + (j & 3u);
        const uint32_t ty = f.ty0 + (lband * f.band_count + f.band_index) * 8u + (j >> 2);
#else                    // four 32x1 tiles per 32x4 macro tile (mcols counts half-height macro tiles: two per 32 columns)
        const uint32_t tx = f.tx0 + (mcol >> 1) * 32u + j;
        const uint32_t ty = f.ty0 + (lband * f.band_count + f.band_index) * 8u + (mcol & 1u) * 4u + k;
#endif
        if (tx >= f.tx0 + f.tw || ty >= f.ty0 + f.th) return;
        int x = (int)tx, y = (int)ty;
        if (f.checkerboard) { // Renderer.cu:186-194
            y *= 2;
            if ((x % 2) == 0) y += 1;
            if (f.frame_number % 2u == 0u) y += 1;
        }
        if ((uint32_t)x >= f.W || (uint32_t)y >= f.H) return;
        pix = (uint32_t)x | ((uint32_t)y << 15) | ((f.crosshair && tx == (f.W >> 1) && ty == (f.H >> 1)) ? (1u << 30) : 0u);
        if (ORTHO) { // every ray runs along fwd (normalised by Raytrace only), from its own origin
            const V3 o = pixel_origin(f, true, x, y);
            lane_setup(w, L, to_brick_units(w, o), normalize3<NORM>(f.fwd), 0u);
            return;
        }
        // Raytrace normalises the (already normalised) direction it is given once more (VolumeRaytracer.cu:363)
        lane_setup(w, L, f.start0, normalize3<NORM>(primary_dir(x, y)), 0u);
    }

    __device__ __forceinline__ void finish(const WorldView& w, Lane& L, V3 pos, uint32_t ncode) {
        const bool hit = L.flags & F_HIT;
        const bool queued = QUEUED ? hit : (hit && !f.debug_view && f.shadows); // the shadow Raytrace of Renderer.cu:102 runs in shadow_rays
        if (STATS) add_stats(stats, 1u, hit, 0u, 0u, queued ? 0u : (uint32_t)L.total, queued ? Work{0u, 0u, 0u} : L.work);
        if (queued) { // staged in shared memory until the tile is complete, see flush()
            const uint32_t tid = threadIdx.x;
            stage_rec[tid] = make_float4(pos.x, pos.y, pos.z, __uint_as_float(pix));
            stage_aux[tid] = (uint32_t)L.total | (ncode << 16);
            if (STATS) { stage_aux[tid + 128] = L.work.C; stage_aux[tid + 256] = L.work.K; stage_aux[tid + 384] = L.work.E; }
            pix |= 0x80000000u;
        } else {
            const int x = (int)(pix & 32767u), y = (int)((pix >> 15) & 32767u);
            if (QUEUED) emit_pixel<NORM, true>(f, ORTHO, bgra, steps_plane, pix, false, pos, 0u, false, L.total, ORTHO ? f.fwd : primary_dir(x, y));
            else emit_pixel<NORM>(f, ORTHO, bgra, steps_plane, pix, hit, pos, ncode, false, L.total, hit ? pos : (ORTHO ? f.fwd : primary_dir(x, y)));
        }
        L.flags = ST_IDLE;
    }
};

// Sun rays.  Set-up (origin, grid-entry clip, first traversal start: six IEEE divisions per ray) and shading (powf,
// tonemap) are the same straight-line work for every ray, so they run as dense one-thread-per-entry kernels around
// the march instead of inside it, where lanes finish and start one by one:
//   shadow_setup  queue entry -> lane record          (dense)
//   shadow_rays   lane record -> steps | hit          (persistent warps, the divergent part)
//   shade_pixels  entry + result -> pixel             (dense)
template <int NORM>
__device__ __forceinline__ V3 sun_direction(const FrameParams& f, V3& once) {
    // calculateColor, Renderer.cu:98-102: shadowRay = normalize(L); Raytrace normalises what it is given again (:363)
    once = normalize3<NORM>(f.light);
    return normalize3<NORM>(once);
}

#ifndef VX_SETUP_THREADS
#define VX_SETUP_THREADS 256
#define VX_SETUP_MIN_BLOCKS 4
#endif
template <int NORM, bool STATS, bool F8>
__global__ void __launch_bounds__(VX_SETUP_THREADS, VX_SETUP_MIN_BLOCKS) shadow_setup(WorldView w_in, FrameParams f, ShadowQueue q) {
    const WorldView w = F8 ? factor8_view(w_in) : w_in;
    const unsigned long long n = *q.count;
    V3 sray;
    const V3 d = sun_direction<NORM>(f, sray);
    const float ff = (float)w.factor;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        const float4 r = q.rec[i];
        const V3 so = mk(fadd(r.x, fmul(sray.x, 0.01f)), fadd(r.y, fmul(sray.y, 0.01f)), fadd(r.z, fmul(sray.z, 0.01f)));
        Lane L;
        lane_setup(w, L, to_brick_units(w, so), d, F_SHADOW);
        L.tx = L.ty = L.tz = 0.0f; L.idx = 0; L.bidx = 0;
        lane_begin<true>(w, L, kMaxSteps);
        // A sun ray starts just off a surface, i.e. inside an occupied brick: its first examination finds that brick, and the
        // march through it (a handful of voxel steps, much the same for the neighbouring pixels that sit next to each other
        // in the queue) is done here, with all lanes busy, and so is the start of the traversal that follows.  shadow_rays
        // picks the ray up from there.
        if (lane_state(L) == ST_FAST && cell_bit(w.ptop, L.idx)) { L.bidx = L.idx; lane_set_state(L, ST_CAND); }
        if (lane_state(L) == ST_CAND) {
            V3 pos; uint32_t ncode;
            lane_cand<false, STATS, true>(w, L, pos, ncode);
            if (lane_state(L) == ST_BEGIN) lane_begin<true>(w, L, kMaxSteps);
        }
        q.lanes[3 * i] = make_uint4(__float_as_uint(L.s.x), __float_as_uint(L.s.y), __float_as_uint(L.s.z), __float_as_uint(L.tx));
        q.lanes[3 * i + 1] = make_uint4(__float_as_uint(L.ty), __float_as_uint(L.tz), L.idx, (uint32_t)L.total);
        q.lanes[3 * i + 2] = make_uint4(L.pc, L.flags, 0u, 0u);
        if (STATS) { q.aux[i + 5 * f.ids] = L.work.C; q.aux[i + 6 * f.ids] = L.work.K; q.aux[i + 7 * f.ids] = L.work.E; }
    }
}

//
// FUSE: the march shades too, and shade_pixels is not launched.  A finished sun ray leaves (queue entry, result) in a
// per-warp list in shared memory; whenever 32 are waiting the whole warp shades one each — the same dense work as
// shade_pixels, but spread over the march.  For a frame that lives in another GPU's memory (the shared frame of the
// multi-GPU partition, vx_shared_open) that spreads the pixel stores over the whole kernel instead of sending every
// rank's pixels over NVLink in one burst at the end of the frame.
constexpr uint32_t kShadeList = 64; // per warp: fewer than 32 waiting + at most 32 finishing in one round

template <int NORM, bool STATS, bool FUSE>
struct ShadowPolicy {
    static constexpr int kRefill = VX_SHADOW_REFILL;
    static constexpr bool kDrain = FUSE, kSun = true;
    static constexpr int kFastQuad = VX_SHADOW_QUAD, kLeaveNum = VX_SHADOW_LEAVE_NUM, kLeaveDen = VX_SHADOW_LEAVE_DEN;
    const FrameParams& f;
    ShadowQueue q;
    uint8_t* __restrict__ bgra;
    int32_t* __restrict__ steps_plane;
    uint2* list;               // FUSE: this warp's kShadeList entries
    unsigned long long n;
    uint32_t slot;             // per lane: queue entry whose sun ray this lane marches
    uint32_t head, waiting;    // FUSE, warp-uniform: the list is entries head .. head + waiting - 1 (mod kShadeList)
    uint32_t done_ballot;      // FUSE: lanes finishing in this round (set by run_lanes before finish())

    __device__ __forceinline__ ShadowPolicy(const FrameParams& fp, ShadowQueue sq, uint8_t* b, int32_t* sp, uint2* l)
        : f(fp), q(sq), bgra(b), steps_plane(sp), list(l), n(*sq.count), slot(0), head(0), waiting(0), done_ballot(0) {}
    __device__ __forceinline__ unsigned long long count() const { return n; }
    __device__ __forceinline__ int max_steps() const { return kMaxSteps; }
    __device__ __forceinline__ void prepare(Lane& L) { // direction and slab inverses are the same for every sun ray
        V3 once;
        L.d = sun_direction<NORM>(f, once);
        L.ivx = fdiv(1.0f, L.d.x == 0.0f ? kFltEps : L.d.x); L.ivy = fdiv(1.0f, L.d.y == 0.0f ? kFltEps : L.d.y);
        L.ivz = fdiv(1.0f, L.d.z == 0.0f ? kFltEps : L.d.z);
    }

    __device__ __forceinline__ void init(const WorldView&, unsigned long long id, Lane& L) {
        slot = (uint32_t)id;
        const uint4 a = q.lanes[3 * id], b = q.lanes[3 * id + 1], c = q.lanes[3 * id + 2];
        L.s = mk(__uint_as_float(a.x), __uint_as_float(a.y), __uint_as_float(a.z));
        L.tx = __uint_as_float(a.w); L.ty = __uint_as_float(b.x); L.tz = __uint_as_float(b.y);
        L.idx = b.z; L.total = (int)b.w; L.pc = c.x; L.flags = c.y;
        L.csteps = 0; L.fsteps = 0; L.fcell = 0; L.work.C = L.work.K = L.work.E = 0;
        if (STATS) { L.work.C = q.aux[id + 5 * f.ids]; L.work.K = q.aux[id + 6 * f.ids]; L.work.E = q.aux[id + 7 * f.ids]; } // shadow_setup's share
    }

    __device__ __forceinline__ void finish(const WorldView&, Lane& L, V3, uint32_t) {
        const uint32_t result = (uint32_t)L.total | ((L.flags & F_HIT) ? 0x80000000u : 0u);
        if (FUSE) {
            const uint32_t rank = (uint32_t)__popc(done_ballot & ((1u << (threadIdx.x & 31u)) - 1u));
            list[(head + waiting + rank) & (kShadeList - 1u)] = make_uint2(slot, result);
        } else {
            q.aux[slot + f.ids] = result;
            if (STATS) { q.aux[slot + 5 * f.ids] = L.work.C; q.aux[slot + 6 * f.ids] = L.work.K; q.aux[slot + 7 * f.ids] = L.work.E; }
        }
        L.flags = ST_IDLE;
    }

    __device__ __forceinline__ void shade_waiting(uint32_t count) {
        __syncwarp();
        if ((threadIdx.x & 31u) < count) {
            const uint2 e = list[(head + (threadIdx.x & 31u)) & (kShadeList - 1u)];
            const float4 r = q.rec[e.x];
            const uint32_t a = q.aux[e.x];
            emit_pixel<NORM>(f, f.ortho_w > 0.0f, bgra, steps_plane, __float_as_uint(r.w), true, mk(r.x, r.y, r.z), a >> 16, (e.y >> 31) != 0u,
                             (int)(a & 0xFFFFu) + (int)(e.y & 0x7FFFFFFFu), mk(0.0f, 0.0f, 0.0f));
        }
        __syncwarp();
        head = (head + count) & (kShadeList - 1u);
        waiting -= count;
    }
    __device__ __forceinline__ void drain(uint32_t done) { // all 32 lanes
        waiting += (uint32_t)__popc(done);
        if (waiting >= 32u) shade_waiting(32u);
    }
    __device__ __forceinline__ void flush() { // the whole warp is idle: whatever waits goes out (at the latest, at the end)
        if (FUSE) { if (waiting >= 32u) shade_waiting(32u); if (waiting) shade_waiting(waiting); }
    }
};

template <int NORM, bool STATS>
__global__ void __launch_bounds__(256) shade_pixels(FrameParams f, uint8_t* __restrict__ bgra, int32_t* __restrict__ steps_plane,
                                                    unsigned long long* __restrict__ stats, ShadowQueue q) {
    const unsigned long long n = *q.count;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        const float4 r = q.rec[i];
        const uint32_t a = q.aux[i], b = q.aux[i + f.ids];
        const bool shadowed = b >> 31;
        const int steps = (int)(a & 0xFFFFu) + (int)(b & 0x7FFFFFFFu);
        if (STATS) {
            Work wk;
            wk.C = q.aux[i + 2 * f.ids] + q.aux[i + 5 * f.ids]; wk.K = q.aux[i + 3 * f.ids] + q.aux[i + 6 * f.ids];
            wk.E = q.aux[i + 4 * f.ids] + q.aux[i + 7 * f.ids];
            add_stats(stats, 0u, 0u, 1u, shadowed, (uint32_t)steps, wk);
        }
        emit_pixel<NORM>(f, f.ortho_w > 0.0f, bgra, steps_plane, __float_as_uint(r.w), true, mk(r.x, r.y, r.z), a >> 16, shadowed, steps, mk(0.0f, 0.0f, 0.0f));
    }
}

// Incoherent secondary rays (BASELINE config 5): the ambient-occlusion sample loop of Renderer.cu:121-165, sample i = 0.
// One thread per queued primary hit; ray order follows the queue, `pixels` says which pixel each ray belongs to.
__device__ __forceinline__ float random_float(uint32_t seed) { return fdiv((float)hash32(seed), (float)0xffffffffu); } // cuda_noise.cuh:66-71 (hash: vx_noise.cuh)

template <int NORM>
__global__ void __launch_bounds__(256) bounce_setup(FrameParams f, ShadowQueue q, float* __restrict__ origins, float* __restrict__ dirs,
                                                    uint32_t* __restrict__ pixels) {
    const unsigned long long n = *q.count;
    for (unsigned long long i = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (unsigned long long)gridDim.x * blockDim.x) {
        const float4 r = q.rec[i];
        const uint32_t pix = __float_as_uint(r.w);
        const int x = (int)(pix & 32767u), y = (int)((pix >> 15) & 32767u);
        int ty = y; // thread coordinates: undo the checkerboard row mapping of Renderer.cu:186-194
        if (f.checkerboard) ty = (y - ((x % 2) == 0 ? 1 : 0) - (f.frame_number % 2u == 0u ? 1 : 0)) / 2;
        const V3 nn = normal_from_code(q.aux[i] >> 16);
        const V3 n3 = mk(-nn.x, -nn.y, -nn.z); // the shading normal (screenDispatch negates what Raytrace returns)
        const int seed = ty * (int)f.W + x;
        const int si = seed + 0 * 1000 + ((int)f.frame_number + 1) * 1000;
        V3 sd = mk(fsub(fmul(random_float((uint32_t)si), 2.0f), 1.0f), fsub(fmul(random_float((uint32_t)(si * 10)), 2.0f), 1.0f),
                   fsub(fmul(random_float((uint32_t)(si * 100)), 2.0f), 1.0f));
        sd = normalize3<NORM>(sd);
        if (dot3(sd, n3) < 0.0f) { // reflect(i, n) = i - 2.0f * n * dot(n, i), helper_math.h:1427-1430
            const float dn = dot3(n3, sd);
            sd = mk(fsub(sd.x, fmul(fmul(2.0f, n3.x), dn)), fsub(sd.y, fmul(fmul(2.0f, n3.y), dn)), fsub(sd.z, fmul(fmul(2.0f, n3.z), dn)));
        }
        origins[3 * i] = fadd(r.x, fmul(n3.x, 0.01f)); origins[3 * i + 1] = fadd(r.y, fmul(n3.y, 0.01f)); origins[3 * i + 2] = fadd(r.z, fmul(n3.z, 0.01f));
        dirs[3 * i] = sd.x; dirs[3 * i + 1] = sd.y; dirs[3 * i + 2] = sd.z;
        pixels[i] = (uint32_t)y * f.W + (uint32_t)x;
    }
}

// SPEC: 8^3 bricks (F8 above) and the first launch of a shadowed frame (PrimaryPolicy's QUEUED).
template <int NORM, bool STATS, bool ORTHO, bool SPEC>
__global__ void __launch_bounds__(128, VX_MIN_BLOCKS) primary_rays(WorldView w_in, FrameParams f, uint8_t* __restrict__ bgra, int32_t* __restrict__ steps_plane,
                                                                   unsigned long long* __restrict__ stats, ShadowQueue q, unsigned long long* next) {
    __shared__ float4 stage_rec[128];
    __shared__ uint32_t stage_aux[STATS ? 512 : 128];
    __shared__ uint32_t tile_slots[8];
    if (threadIdx.x < 8) tile_slots[threadIdx.x] = 0xFFFFFFFFu;
    __syncthreads();
    const WorldView w = SPEC ? factor8_view(w_in) : w_in;
    PrimaryPolicy<NORM, STATS, ORTHO, SPEC> pol(f, bgra, steps_plane, stats, q);
    pol.stage_rec = stage_rec; pol.stage_aux = stage_aux; pol.tile_slot = tile_slots + 2 * (threadIdx.x >> 5);
    run_lanes<false, STATS>(w, pol, next);
}
template <int NORM, bool STATS, bool FUSE, bool F8>
__global__ void __launch_bounds__(128, VX_SHADOW_MIN_BLOCKS) shadow_rays(WorldView w_in, FrameParams f, uint8_t* __restrict__ bgra, int32_t* __restrict__ steps_plane,
                                                                  ShadowQueue q, unsigned long long* next) {
    __shared__ uint2 shade_list[FUSE ? 4 * kShadeList : 1];
    const WorldView w = F8 ? factor8_view(w_in) : w_in;
    ShadowPolicy<NORM, STATS, FUSE> pol(f, q, bgra, steps_plane, shade_list + (FUSE ? (threadIdx.x >> 5) * kShadeList : 0u));
    run_lanes<false, STATS>(w, pol, next);
}

__global__ void tanf_kernel(float x, float* out) { *out = tanf(x); }
__global__ void normalize_kernel(const float* __restrict__ in, float* __restrict__ out, size_t n) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const V3 v = normalize3<0>(mk(in[3 * i], in[3 * i + 1], in[3 * i + 2]));
    out[3 * i] = v.x; out[3 * i + 1] = v.y; out[3 * i + 2] = v.z;
}

} // namespace vx

// =============================================================================================
// host side
// =============================================================================================
using namespace vx;

extern "C" int vx_tracer_create(vx_tracer** out, vx_world* w, size_t max_rays) {
    VX_REQUIRE(out && w, "NULL argument");
    *out = nullptr;
    DeviceGuard guard(w->device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    vx_tracer* t = new vx_tracer();
    t->world = w; t->device = w->device; t->max_rays = max_rays;
    const size_t n = max_rays ? max_rays : 1;
#define VX_ALLOC(ptr, bytes) do { cudaError_t e_ = cudaMalloc((void**)&(ptr), (bytes)); if (e_ != cudaSuccess) { vx_tracer_destroy(t); return cuda_fail(e_, "cudaMalloc", __FILE__, __LINE__); } } while (0)
    VX_ALLOC(t->d_origins, n * 12); VX_ALLOC(t->d_dirs, n * 12);
    VX_ALLOC(t->d_hit, n); VX_ALLOC(t->d_pos, n * 12); VX_ALLOC(t->d_normal, n * 12); VX_ALLOC(t->d_steps, n * 4);
    VX_ALLOC(t->d_voxel, n * 12); VX_ALLOC(t->d_brick, n * 4); VX_ALLOC(t->d_csteps, n * 4); VX_ALLOC(t->d_fsteps, n * 4);
    VX_ALLOC(t->d_counters, n * 12);
    VX_ALLOC(t->d_next, kNextSlots * sizeof(unsigned long long));
#undef VX_ALLOC
    cudaDeviceGetAttribute(&t->sm_count, cudaDevAttrMultiProcessorCount, t->device);
    VX_CUDA(cudaStreamCreateWithFlags(&t->stream, cudaStreamNonBlocking));
    *out = t;
    return VX_OK;
}

extern "C" int vx_tracer_destroy(vx_tracer* t) {
    if (!t) return VX_OK;
    DeviceGuard guard(t->device);
    cudaFree(t->d_origins); cudaFree(t->d_dirs); cudaFree(t->d_hit); cudaFree(t->d_pos); cudaFree(t->d_normal);
    cudaFree(t->d_steps); cudaFree(t->d_voxel); cudaFree(t->d_brick); cudaFree(t->d_csteps); cudaFree(t->d_fsteps);
    cudaFree(t->d_counters); cudaFree(t->d_frame); cudaFree(t->d_next); cudaFree(t->d_queue_rec); cudaFree(t->d_queue_aux); cudaFree(t->d_queue_lanes);
    for (int i = 0; i < 5; ++i) if (t->ev[i]) cudaEventDestroy(t->ev[i]);
    for (int i = 0; i < 4; ++i) if (t->slice_ev[i]) cudaEventDestroy(t->slice_ev[i]);
    if (t->copy_stream) cudaStreamDestroy(t->copy_stream);
    if (t->stream) cudaStreamDestroy(t->stream);
    destroy_batch_pipe(t->batch);
    cudaFree(t->d_tile_trace);
    delete t;
    return VX_OK;
}

// Persistent launch: as many 128-thread blocks as stay resident (or fewer when there is less work), plus a fresh
// zeroed ray counter taken from a small ring so launches on different streams do not share one.
template <class K>
static int persistent_blocks(vx_tracer* t, K kernel, unsigned long long ids, unsigned& blocks) {
    int per_sm = 0;
    VX_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, 128, 0));
    if (per_sm < 1) per_sm = 1;
    static const int cap = [] { const char* e = getenv("VX_BLOCKS_PER_SM"); return e ? atoi(e) : 0; }(); // experiments: fewer resident blocks
    if (cap > 0 && cap < per_sm) per_sm = cap;
    const unsigned long long full = (unsigned long long)per_sm * (unsigned long long)(t->sm_count > 0 ? t->sm_count : 148);
    const unsigned long long need = (ids + 127) / 128;
    blocks = (unsigned)(need < full ? need : full);
    if (blocks == 0) blocks = 1;
    return VX_OK;
}
static int next_counter(vx_tracer* t, cudaStream_t s, unsigned long long** out) {
    unsigned long long* p = t->d_next + (t->next_seq++ % kNextSlots);
    VX_CUDA(cudaMemsetAsync(p, 0, sizeof(unsigned long long), s));
    *out = p;
    return VX_OK;
}

// The traversal grid (one occupancy bit per coarse cell, 8.7 MB at 8192x512x8192) is what every ray of every frame reads,
// cell by cell; bricks and records are touched once or twice.  An access-policy window marks it persisting in L2 for the
// kernels launched on `s`, so that the streaming traffic of a frame (records, brick masks, the frame queues, the pixels)
// and whatever else runs between frames does not evict it.  Best effort: a device without persisting L2 just runs without.
// VX_L2_PIN=0 switches it off.
static void pin_top_level(vx_tracer* t, cudaStream_t s) {
    static const bool enabled = [] { const char* e = getenv("VX_L2_PIN"); return !(e && atoi(e) == 0); }();
    const vx_world* w = t->world;
    if (!enabled || !w || !w->d_ptop) return;
    if (t->l2_tried && t->l2_stream == s && t->l2_base == w->d_ptop) return;
    t->l2_tried = true; t->l2_stream = s; t->l2_base = w->d_ptop;
    const size_t pcells = (size_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
    size_t bytes = (pcells + 31) / 32 * 4;
    int max_window = 0, max_persist = 0;
    cudaDeviceGetAttribute(&max_window, cudaDevAttrMaxAccessPolicyWindowSize, t->device);
    cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, t->device);
    if (max_window <= 0 || max_persist <= 0) { cudaGetLastError(); return; }
    if (bytes > (size_t)max_window) bytes = (size_t)max_window;
    size_t carve = 0;
    cudaDeviceGetLimit(&carve, cudaLimitPersistingL2CacheSize);
    const size_t want = bytes < (size_t)max_persist ? bytes : (size_t)max_persist;
    if (carve < want) cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, want);
    cudaStreamAttrValue attr = {};
    attr.accessPolicyWindow.base_ptr = (void*)w->d_ptop;
    attr.accessPolicyWindow.num_bytes = bytes;
    attr.accessPolicyWindow.hitRatio = want >= bytes ? 1.0f : (float)((double)want / (double)bytes);
    attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
    attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
    cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr);
    cudaGetLastError(); // best effort
}

static int launch_trace(vx_tracer* t, const float* d_o, const float* d_d, size_t n, const vx_trace_opts* opts, const vx_hits* out,
                        cudaStream_t s) {
    if (n == 0) return VX_OK;
    const int norm = opts ? opts->norm_mode : VX_NORM_DEVICE;
    const int max_steps = (opts && opts->max_steps > 0) ? opts->max_steps : kMaxSteps;
    VX_REQUIRE(norm >= 0 && norm <= 2, "bad norm_mode");
    HitPlanes p{out->hit, out->pos, out->normal, out->steps, out->voxel, out->brick, out->coarse_steps, out->fine_steps, out->counters};
    const WorldView w = t->world->view();
    pin_top_level(t, s);
    const bool count = out->counters != nullptr;
    unsigned long long* next = nullptr;
    int rc = next_counter(t, s, &next);
    if (rc) return rc;
    unsigned blocks = 1;
#define VX_LAUNCH(N, C, F) do { rc = persistent_blocks(t, trace_rays<N, C, F>, n, blocks); if (rc) return rc; \
        trace_rays<N, C, F><<<blocks, 128, 0, s>>>(w, d_o, d_d, (unsigned long long)n, max_steps, p, next); } while (0)
    if (count) { if (norm == 0) VX_LAUNCH(0, true, false); else if (norm == 1) VX_LAUNCH(1, true, false); else VX_LAUNCH(2, true, false); }
    else if (w.factor == 8) { if (norm == 0) VX_LAUNCH(0, false, true); else if (norm == 1) VX_LAUNCH(1, false, true); else VX_LAUNCH(2, false, true); }
    else { if (norm == 0) VX_LAUNCH(0, false, false); else if (norm == 1) VX_LAUNCH(1, false, false); else VX_LAUNCH(2, false, false); }
#undef VX_LAUNCH
    VX_CUDA(cudaGetLastError());
    return VX_OK;
}

extern "C" int vx_trace_device(vx_tracer* t, const float* d_origins, const float* d_dirs, size_t n, const vx_trace_opts* opts,
                               const vx_hits* out_dev, void* stream) {
    VX_REQUIRE(t && out_dev && (n == 0 || (d_origins && d_dirs)), "NULL argument");
    DeviceGuard guard(t->device);
    return launch_trace(t, d_origins, d_dirs, n, opts, out_dev, (cudaStream_t)stream);
}

// ---- batch rays in HOST arrays (VoxelRaytracer3D::Raytrace(vector, vector), VolumeRaytracer.cu:574-618) ----------------------
// The reference copies both vectors up with blocking pageable cudaMemcpy, runs the kernel, and copies three planes down the
// same way.  A pageable copy is a single host thread feeding the driver's staging buffers (7-10 GB/s), which for 2 M rays is
// 20 ms around a 0.2 ms march.  Here the call is a three-stage pipeline over chunks of kBatchChunk rays: a small pool of
// host threads copies chunk k+1's rays into a page-locked ring while the DMA engines move chunk k up and chunk k-1's planes
// down (own streams, events between them) and the march of chunk k runs in between; only the planes the caller passed are
// computed, moved and copied out.  Ring, pool and streams belong to the tracer and are created on the first call.
namespace vx {

constexpr size_t kBatchChunk = 256 * 1024; // rays per pipeline stage
constexpr int kBatchSlots = 3;             // ring depth: staging in, in flight, draining

struct CopyPool { // fixed set of worker threads running one sliced memcpy job at a time
    std::vector<std::thread> workers;
    std::mutex mu;
    std::condition_variable cv_job, cv_done;
    struct Part { void* dst; const void* src; size_t bytes; };
    std::vector<Part> parts;
    size_t next_part = 0, pending = 0;
    unsigned long long generation = 0;
    bool stop = false;

    explicit CopyPool(int n) {
        for (int i = 0; i < n; ++i) workers.emplace_back([this] { run(); });
    }
    ~CopyPool() {
        { std::lock_guard<std::mutex> g(mu); stop = true; }
        cv_job.notify_all();
        for (auto& w : workers) w.join();
    }
    void run() {
        std::unique_lock<std::mutex> lk(mu);
        for (;;) {
            cv_job.wait(lk, [this] { return stop || next_part < parts.size(); });
            if (stop) return;
            const Part p = parts[next_part++];
            lk.unlock();
            memcpy(p.dst, p.src, p.bytes);
            lk.lock();
            if (--pending == 0) cv_done.notify_all();
        }
    }
    // copies every (dst, src, bytes) of `jobs`, each cut into slices, on the workers and the calling thread; returns when done
    void copy(const std::vector<Part>& jobs) {
        constexpr size_t kSlice = 256 * 1024;
        std::unique_lock<std::mutex> lk(mu);
        parts.clear(); next_part = 0;
        for (const Part& j : jobs)
            for (size_t off = 0; off < j.bytes; off += kSlice)
                parts.push_back({(char*)j.dst + off, (const char*)j.src + off, j.bytes - off < kSlice ? j.bytes - off : kSlice});
        pending = parts.size();
        if (!pending) return;
        cv_job.notify_all();
        while (next_part < parts.size()) { // the caller works too
            const Part p = parts[next_part++];
            lk.unlock();
            memcpy(p.dst, p.src, p.bytes);
            lk.lock();
            --pending;
        }
        cv_done.wait(lk, [this] { return pending == 0; });
    }
};

struct BatchPipe {
    CopyPool* pool = nullptr;
    uint8_t* h_in[kBatchSlots] = {};   // pinned: origins then dirs of one chunk
    uint8_t* h_out[kBatchSlots] = {};  // pinned: the requested planes of one chunk, back to back
    cudaStream_t s_up = nullptr, s_run = nullptr, s_down = nullptr;
    cudaEvent_t ev_up[kBatchSlots] = {}, ev_run[kBatchSlots] = {}, ev_down[kBatchSlots] = {};
    ~BatchPipe() {
        delete pool;
        for (int i = 0; i < kBatchSlots; ++i) {
            if (h_in[i]) cudaFreeHost(h_in[i]);
            if (h_out[i]) cudaFreeHost(h_out[i]);
            if (ev_up[i]) cudaEventDestroy(ev_up[i]);
            if (ev_run[i]) cudaEventDestroy(ev_run[i]);
            if (ev_down[i]) cudaEventDestroy(ev_down[i]);
        }
        if (s_up) cudaStreamDestroy(s_up);
        if (s_run) cudaStreamDestroy(s_run);
        if (s_down) cudaStreamDestroy(s_down);
    }
};
void destroy_batch_pipe(BatchPipe* p) { delete p; }

constexpr size_t kPlaneBytes[9] = {1, 12, 12, 4, 12, 4, 4, 4, 12}; // hit pos normal steps voxel brick coarse fine counters
constexpr size_t kOutBytesPerRay = 1 + 12 + 12 + 4 + 12 + 4 + 4 + 4 + 12;

static int batch_pipe(vx_tracer* t, BatchPipe** out) {
    if (t->batch) { *out = t->batch; return VX_OK; }
    BatchPipe* p = new BatchPipe();
    t->batch = p; // owned by the tracer from here on (vx_tracer_destroy frees whatever was created)
    static const int kThreads = [] {
        const char* e = getenv("VX_BATCH_THREADS");
        int v = e ? atoi(e) : (int)std::thread::hardware_concurrency() / 2;
        return v < 1 ? 1 : (v > 16 ? 16 : v);
    }();
    p->pool = new CopyPool(kThreads - 1);
    for (int i = 0; i < kBatchSlots; ++i) {
        VX_CUDA(cudaHostAlloc((void**)&p->h_in[i], kBatchChunk * 24, cudaHostAllocDefault));
        VX_CUDA(cudaHostAlloc((void**)&p->h_out[i], kBatchChunk * kOutBytesPerRay, cudaHostAllocDefault));
        VX_CUDA(cudaEventCreateWithFlags(&p->ev_up[i], cudaEventDisableTiming));
        VX_CUDA(cudaEventCreateWithFlags(&p->ev_run[i], cudaEventDisableTiming));
        VX_CUDA(cudaEventCreateWithFlags(&p->ev_down[i], cudaEventDisableTiming));
    }
    VX_CUDA(cudaStreamCreateWithFlags(&p->s_up, cudaStreamNonBlocking));
    VX_CUDA(cudaStreamCreateWithFlags(&p->s_run, cudaStreamNonBlocking));
    VX_CUDA(cudaStreamCreateWithFlags(&p->s_down, cudaStreamNonBlocking));
    *out = p;
    return VX_OK;
}

// A handful of rays (probes, unit tests): not worth a ring — plain copies on the tracer's stream.
constexpr size_t kBatchSmall = 32 * 1024;
static int trace_small(vx_tracer* t, const float* origins, const float* dirs, size_t n, const vx_trace_opts* opts, const vx_hits* out) {
    cudaStream_t s = t->stream;
    VX_CUDA(cudaMemcpyAsync(t->d_origins, origins, n * 12, cudaMemcpyHostToDevice, s));
    VX_CUDA(cudaMemcpyAsync(t->d_dirs, dirs, n * 12, cudaMemcpyHostToDevice, s));
    void* const h_plane[9] = {out->hit, out->pos, out->normal, out->steps, out->voxel, out->brick, out->coarse_steps, out->fine_steps, out->counters};
    void* const d_plane[9] = {t->d_hit, t->d_pos, t->d_normal, t->d_steps, t->d_voxel, t->d_brick, t->d_csteps, t->d_fsteps, t->d_counters};
    vx_hits dev;
    dev.hit = out->hit ? t->d_hit : nullptr; dev.pos = out->pos ? t->d_pos : nullptr; dev.normal = out->normal ? t->d_normal : nullptr;
    dev.steps = out->steps ? t->d_steps : nullptr; dev.voxel = out->voxel ? t->d_voxel : nullptr; dev.brick = out->brick ? t->d_brick : nullptr;
    dev.coarse_steps = out->coarse_steps ? t->d_csteps : nullptr; dev.fine_steps = out->fine_steps ? t->d_fsteps : nullptr;
    dev.counters = out->counters ? t->d_counters : nullptr;
    int rc = launch_trace(t, t->d_origins, t->d_dirs, n, opts, &dev, s);
    if (rc) return rc;
    for (int i = 0; i < 9; ++i)
        if (h_plane[i]) VX_CUDA(cudaMemcpyAsync(h_plane[i], d_plane[i], n * kPlaneBytes[i], cudaMemcpyDeviceToHost, s));
    VX_CUDA(cudaStreamSynchronize(s));
    return VX_OK;
}

} // namespace vx

extern "C" int vx_trace(vx_tracer* t, const float* origins, const float* dirs, size_t n, const vx_trace_opts* opts, const vx_hits* out) {
    VX_REQUIRE(t && out && (n == 0 || (origins && dirs)), "NULL argument");
    if (n > t->max_rays) { set_error("vx_trace: more rays than the tracer was created for"); return VX_ERR_CAPACITY; }
    if (n == 0) return VX_OK;
    DeviceGuard guard(t->device);
    if (n < kBatchSmall) return trace_small(t, origins, dirs, n, opts, out);
    BatchPipe* p = nullptr;
    int rc = batch_pipe(t, &p);
    if (rc) return rc;
    void* const h_plane[9] = {out->hit, out->pos, out->normal, out->steps, out->voxel, out->brick, out->coarse_steps, out->fine_steps, out->counters};
    void* const d_plane[9] = {t->d_hit, t->d_pos, t->d_normal, t->d_steps, t->d_voxel, t->d_brick, t->d_csteps, t->d_fsteps, t->d_counters};
    const size_t chunks = (n + kBatchChunk - 1) / kBatchChunk;
    auto count_of = [&](size_t k) { return k + 1 < chunks ? kBatchChunk : n - k * kBatchChunk; };
    // drain(k): chunk k's planes from the ring into the caller's arrays
    auto drain = [&](size_t k) -> int {
        const int slot = (int)(k % kBatchSlots);
        const size_t first = k * kBatchChunk, cnt = count_of(k);
        VX_CUDA(cudaEventSynchronize(p->ev_down[slot]));
        std::vector<CopyPool::Part> jobs;
        size_t off = 0;
        for (int i = 0; i < 9; ++i)
            if (h_plane[i]) { jobs.push_back({(char*)h_plane[i] + first * kPlaneBytes[i], p->h_out[slot] + off, cnt * kPlaneBytes[i]}); off += cnt * kPlaneBytes[i]; }
        p->pool->copy(jobs);
        return VX_OK;
    };
    for (size_t k = 0; k < chunks; ++k) {
        const int slot = (int)(k % kBatchSlots);
        const size_t first = k * kBatchChunk, cnt = count_of(k);
        if (k >= (size_t)kBatchSlots) { rc = drain(k - kBatchSlots); if (rc) return rc; } // frees this slot's halves of the ring
        p->pool->copy({{p->h_in[slot], origins + 3 * first, cnt * 12}, {p->h_in[slot] + cnt * 12, dirs + 3 * first, cnt * 12}});
        VX_CUDA(cudaMemcpyAsync(t->d_origins + 3 * first, p->h_in[slot], cnt * 12, cudaMemcpyHostToDevice, p->s_up));
        VX_CUDA(cudaMemcpyAsync(t->d_dirs + 3 * first, p->h_in[slot] + cnt * 12, cnt * 12, cudaMemcpyHostToDevice, p->s_up));
        VX_CUDA(cudaEventRecord(p->ev_up[slot], p->s_up));
        VX_CUDA(cudaStreamWaitEvent(p->s_run, p->ev_up[slot], 0));
        vx_hits dev;
        dev.hit = out->hit ? t->d_hit + first : nullptr; dev.pos = out->pos ? t->d_pos + 3 * first : nullptr;
        dev.normal = out->normal ? t->d_normal + 3 * first : nullptr; dev.steps = out->steps ? t->d_steps + first : nullptr;
        dev.voxel = out->voxel ? t->d_voxel + 3 * first : nullptr; dev.brick = out->brick ? t->d_brick + first : nullptr;
        dev.coarse_steps = out->coarse_steps ? t->d_csteps + first : nullptr; dev.fine_steps = out->fine_steps ? t->d_fsteps + first : nullptr;
        dev.counters = out->counters ? t->d_counters + 3 * first : nullptr;
        rc = launch_trace(t, t->d_origins + 3 * first, t->d_dirs + 3 * first, cnt, opts, &dev, p->s_run);
        if (rc) return rc;
        VX_CUDA(cudaEventRecord(p->ev_run[slot], p->s_run));
        VX_CUDA(cudaStreamWaitEvent(p->s_down, p->ev_run[slot], 0));
        size_t off = 0;
        for (int i = 0; i < 9; ++i)
            if (h_plane[i]) {
                VX_CUDA(cudaMemcpyAsync(p->h_out[slot] + off, (const char*)d_plane[i] + first * kPlaneBytes[i], cnt * kPlaneBytes[i], cudaMemcpyDeviceToHost, p->s_down));
                off += cnt * kPlaneBytes[i];
            }
        VX_CUDA(cudaEventRecord(p->ev_down[slot], p->s_down));
    }
    for (size_t k = chunks > (size_t)kBatchSlots ? chunks - kBatchSlots : 0; k < chunks; ++k) { rc = drain(k); if (rc) return rc; }
    return VX_OK;
}

extern "C" int vx_hits_postprocess(const float* origins, const float* pos, size_t n, uint8_t* valid, float* distance) {
    // VolumeRaytracer.cu:601-614; distance is computed for misses too (inf), as the reference does
    VX_REQUIRE(origins && pos, "NULL argument");
    for (size_t i = 0; i < n; ++i) {
        const float inf = __builtin_inff();
        if (valid) valid[i] = pos[3 * i] != inf && pos[3 * i + 1] != inf && pos[3 * i + 2] != inf;
        if (distance) {
            const float dx = origins[3 * i] - pos[3 * i], dy = origins[3 * i + 1] - pos[3 * i + 1], dz = origins[3 * i + 2] - pos[3 * i + 2];
            distance[i] = __builtin_sqrtf(dx * dx + dy * dy + dz * dz);
        }
    }
    return VX_OK;
}

extern "C" void vx_get_directions(const float e[3], float fwd[3], float up[3], float right[3]) {
    // Graphics::GetDirections, Renderer.cu:27-42 (host path: float overloads of cos/sin)
    const float cxr = __builtin_cosf(e[0]), sxr = __builtin_sinf(e[0]), cyr = __builtin_cosf(e[1]), syr = __builtin_sinf(e[1]);
    const float fx = cxr * syr, fy = -sxr, fz = cxr * cyr;
    const float rx = cyr, ry = 0.0f, rz = -syr;
    const float ux = fy * rz - fz * ry, uy = fz * rx - fx * rz, uz = fx * ry - fy * rx;
    fwd[0] = fx * -1.0f; fwd[1] = fy * -1.0f; fwd[2] = fz * -1.0f;
    up[0] = ux * -1.0f; up[1] = uy * -1.0f; up[2] = uz * -1.0f;
    right[0] = rx; right[1] = ry; right[2] = rz;
}

extern "C" int vx_device_tanf(float x, float* out, int device) {
    VX_REQUIRE(out != nullptr, "out is NULL");
    DeviceGuard guard(device);
    float* d = nullptr;
    VX_CUDA(cudaMalloc(&d, 4));
    tanf_kernel<<<1, 1>>>(x, d);
    cudaError_t e = cudaMemcpy(out, d, 4, cudaMemcpyDeviceToHost);
    cudaFree(d);
    VX_CUDA(e);
    return VX_OK;
}

extern "C" int vx_device_normalize(const float* dirs, float* out, size_t n, int device) {
    VX_REQUIRE(dirs && out, "NULL argument");
    if (!n) return VX_OK;
    DeviceGuard guard(device);
    float *di = nullptr, *d_out = nullptr;
    VX_CUDA(cudaMalloc(&di, n * 12));
    VX_CUDA(cudaMalloc(&d_out, n * 12));
    VX_CUDA(cudaMemcpy(di, dirs, n * 12, cudaMemcpyHostToDevice));
    normalize_kernel<<<(unsigned)((n + 255) / 256), 256>>>(di, d_out, n);
    cudaError_t e = cudaMemcpy(out, d_out, n * 12, cudaMemcpyDeviceToHost);
    cudaFree(di); cudaFree(d_out);
    VX_CUDA(e);
    return VX_OK;
}

// The launches of one frame (see ShadowQueue above) on stream `cs`, shared by vx_render, vx_bounce_rays' primary pass and
// vx_render_static.  `queue_pass`: the launch runs the sun-ray kernels too; `q` non-null means the primaries queue their hits.
// Kernel variants: STATS when the caller counts; otherwise, for a world of 8^3 bricks, the kernels compiled for that factor
// (and, for the primaries of a queued frame, without shading code).  Optional per-kernel timing (vx_tracer_set_timing):
// events between the launches.
template <int N, bool S, bool F8>
static int launch_frame_as(vx_tracer* t, const WorldView& w, const FrameParams& f, void* d_bgra, int32_t* d_steps, unsigned long long* st,
                           const ShadowQueue& q, unsigned long long* next1, unsigned long long* next2, bool sun_pass, bool fuse, cudaStream_t cs) {
    int rc = VX_OK;
    unsigned blocks = 1;
    const unsigned dense_blocks = (unsigned)(t->sm_count > 0 ? t->sm_count : 148) * 8u; // grid-stride over a device-side count
    const bool queued = q.count != nullptr; // shadows on, debug view off
#define VX_MARK(i) do { if (t->timing) VX_CUDA(cudaEventRecord(t->ev[i], cs)); } while (0)
#define VX_PRIMARY(ORTHO, SPEC) do { \
        rc = persistent_blocks(t, primary_rays<N, S, ORTHO, SPEC>, f.ids, blocks); if (rc) return rc; \
        VX_MARK(0); \
        primary_rays<N, S, ORTHO, SPEC><<<blocks, 128, 0, cs>>>(w, f, (uint8_t*)d_bgra, d_steps, st, q, next1); } while (0)
    t->timed_kernels = 0;
    if (f.ortho_w > 0.0f) { if (F8 && queued) VX_PRIMARY(true, F8); else VX_PRIMARY(true, false); }
    else { if (F8 && queued) VX_PRIMARY(false, F8); else VX_PRIMARY(false, false); }
    VX_MARK(1); t->timed_kernels = 1;
    if (sun_pass) {
        shadow_setup<N, S, F8><<<dense_blocks * (256 / VX_SETUP_THREADS), VX_SETUP_THREADS, 0, cs>>>(w, f, q);
        VX_MARK(2);
        if (!S && fuse) {
            rc = persistent_blocks(t, shadow_rays<N, false, true, F8>, f.ids, blocks); if (rc) return rc;
            shadow_rays<N, false, true, F8><<<blocks, 128, 0, cs>>>(w, f, (uint8_t*)d_bgra, d_steps, q, next2);
            VX_MARK(3);
        } else {
            rc = persistent_blocks(t, shadow_rays<N, S, false, F8>, f.ids, blocks); if (rc) return rc;
            shadow_rays<N, S, false, F8><<<blocks, 128, 0, cs>>>(w, f, (uint8_t*)d_bgra, d_steps, q, next2);
            VX_MARK(3);
            shade_pixels<N, S><<<dense_blocks, 256, 0, cs>>>(f, (uint8_t*)d_bgra, d_steps, st, q);
        }
        VX_MARK(4); t->timed_kernels = 4;
    }
#undef VX_PRIMARY
#undef VX_MARK
    VX_CUDA(cudaGetLastError());
    return VX_OK;
}
static int launch_frame(vx_tracer* t, const WorldView& w, const FrameParams& f, void* d_bgra, int32_t* d_steps, unsigned long long* st,
                        const ShadowQueue& q, unsigned long long* next1, unsigned long long* next2, bool norm_host, bool sun_pass, bool fuse,
                        cudaStream_t cs) {
#define VX_GO(N, S, F8) return launch_frame_as<N, S, F8>(t, w, f, d_bgra, d_steps, st, q, next1, next2, sun_pass, fuse, cs)
    const bool f8 = w.factor == 8;
    if (norm_host) { if (st) VX_GO(1, true, false); else if (f8) VX_GO(1, false, true); else VX_GO(1, false, false); }
    else { if (st) VX_GO(0, true, false); else if (f8) VX_GO(0, false, true); else VX_GO(0, false, false); }
#undef VX_GO
}

static int fill_frame(vx_tracer* t, const vx_camera* cam, const vx_env* env, const vx_render_opts* o, uint32_t W, uint32_t H, FrameParams& f) {
    VX_REQUIRE(W > 0 && H > 0, "empty frame");
    VX_REQUIRE(W < 32768 && H < 32768, "frame dimensions must be below 32768");
    f.origin = V3{cam->origin[0], cam->origin[1], cam->origin[2]};
    const float ff = (float)t->world->factor;
    f.start0 = V3{cam->origin[0] / ff, cam->origin[1] / ff, cam->origin[2] / ff}; // start = origin / factor, VolumeRaytracer.cu:361
    const bool ortho = cam->ortho_w > 0.0f && cam->ortho_h > 0.0f;
    f.ortho_w = ortho ? cam->ortho_w : 0.0f; f.ortho_h = ortho ? cam->ortho_h : 0.0f; f.factor_f = ff;
    f.fwd = V3{cam->fwd[0], cam->fwd[1], cam->fwd[2]};
    f.up = V3{cam->up[0], cam->up[1], cam->up[2]};
    f.right = V3{cam->right[0], cam->right[1], cam->right[2]};
    f.light = V3{env->light_dir[0], env->light_dir[1], env->light_dir[2]};
    f.light_color = V3{env->light_color[0], env->light_color[1], env->light_color[2]};
    f.ambient = V3{env->ambient[0], env->ambient[1], env->ambient[2]};
    if (cam->tan_half > 0.0f) f.tan_half = cam->tan_half;
    else {
        // float fov = FOV * 3.1415 / 180.0; tanf(fov / 2.0f) — Renderer.cu:50-52, evaluated on the device once
        const float fov = (float)((double)cam->fov_deg * 3.1415 / 180.0);
        const float arg = fov / 2.0f;
        if (t->tan_arg != arg) {
            int rc = vx_device_tanf(arg, &t->tan_val, t->device);
            if (rc) return rc;
            t->tan_arg = arg;
        }
        f.tan_half = t->tan_val;
    }
    f.W = W; f.H = H;
    f.frame_number = o->frame_number;
    f.checkerboard = o->checkerboard; f.debug_view = o->debug_view; f.shadows = o->shadows; f.crosshair = o->crosshair;
    const uint32_t rows = o->checkerboard ? (H >> 1) : H; // RenderScreen, Renderer.cu:311-316
    const uint32_t cols = (W + 31u) / 32u * 32u;
    if (o->tile_w && o->tile_h) {
        f.tx0 = o->tile_x0; f.ty0 = o->tile_y0;
        f.tw = o->tile_x0 < cols ? (o->tile_w < cols - o->tile_x0 ? o->tile_w : cols - o->tile_x0) : 0;
        f.th = o->tile_y0 < rows ? (o->tile_h < rows - o->tile_y0 ? o->tile_h : rows - o->tile_y0) : 0;
    } else { f.tx0 = 0; f.ty0 = 0; f.tw = cols; f.th = rows; }
    f.band_count = o->band_count ? o->band_count : 1u;
    f.band_index = o->band_count ? o->band_index : 0u;
    VX_REQUIRE(f.band_index < f.band_count, "band_index must be below band_count");
    return VX_OK;
}

extern "C" int vx_render(vx_tracer* t, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts, void* d_bgra, int32_t* d_steps,
                         uint32_t W, uint32_t H, void* stream) {
    VX_REQUIRE(t && cam && env && opts && d_bgra, "NULL argument");
    VX_REQUIRE(opts->norm_mode == VX_NORM_DEVICE || opts->norm_mode == VX_NORM_HOST, "render supports VX_NORM_DEVICE or VX_NORM_HOST");
    DeviceGuard guard(t->device);
    FrameParams f;
    int rc = fill_frame(t, cam, env, opts, W, H, f);
    if (rc) return rc;
    if (f.tw == 0 || f.th == 0) return VX_OK;
    const uint32_t bands = (f.th + 7u) / 8u;
    const uint32_t my_bands = bands > f.band_index ? (bands - f.band_index + f.band_count - 1u) / f.band_count : 0u;
    if (my_bands == 0) return VX_OK;
#if VX_TILE_SHAPE == 2
    f.mcols = (f.tw + 31u) / 32u * 2u;
#else
    f.mcols = (f.tw + 15u) / 16u;
#endif
    f.ids = (unsigned long long)f.mcols * my_bands * 128ull;
    VX_REQUIRE(f.ids < (1ull << 32), "frame too large for one launch");
    const WorldView w = t->world->view();
    unsigned long long* st = (unsigned long long*)opts->d_stats;
    cudaStream_t cs = (cudaStream_t)stream;
    pin_top_level(t, cs);
    f.tile_trace = nullptr;
    static const bool trace = [] { const char* e = getenv("VX_TILE_TRACE"); return e && atoi(e) != 0; }();
    if (trace) { // diagnostics: a timeline of the launch's tiles (tools/tile_trace.py)
        const size_t words = (size_t)(f.ids >> 5) * 2;
        if (t->tile_trace_words < words) {
            VX_CUDA(cudaStreamSynchronize(cs));
            cudaFree(t->d_tile_trace); t->d_tile_trace = nullptr; t->tile_trace_words = 0;
            VX_CUDA(cudaMalloc(&t->d_tile_trace, words * 4));
            t->tile_trace_words = words;
        }
        t->tile_trace_tiles = (size_t)(f.ids >> 5);
        f.tile_trace = t->d_tile_trace;
    }
    const bool shadow_pass = opts->shadows && !opts->debug_view;
    ShadowQueue q{nullptr, nullptr, nullptr, nullptr};
    if (shadow_pass) { // per-frame queue of primary hits waiting for their sun ray (one entry per pixel of this launch at most)
        const size_t aux_words = (size_t)f.ids * (st ? 8u : 2u);
        if (t->queue_cap < f.ids || t->queue_aux_words < aux_words) {
            VX_CUDA(cudaStreamSynchronize(cs)); // frames in flight on this stream may still read the old queue
            cudaFree(t->d_queue_rec); cudaFree(t->d_queue_aux); cudaFree(t->d_queue_lanes);
            t->d_queue_rec = nullptr; t->d_queue_aux = nullptr; t->d_queue_lanes = nullptr; t->queue_cap = 0; t->queue_aux_words = 0;
            VX_CUDA(cudaMalloc(&t->d_queue_rec, (size_t)f.ids * sizeof(float4)));
            VX_CUDA(cudaMalloc(&t->d_queue_aux, aux_words * sizeof(uint32_t)));
            VX_CUDA(cudaMalloc(&t->d_queue_lanes, (size_t)f.ids * 3u * sizeof(uint4)));
            t->queue_cap = f.ids; t->queue_aux_words = aux_words;
        }
        q.rec = (float4*)t->d_queue_rec; q.aux = (uint32_t*)t->d_queue_aux; q.lanes = (uint4*)t->d_queue_lanes;
        unsigned long long* qc = nullptr;
        rc = next_counter(t, cs, &qc);
        if (rc) return rc;
        q.count = qc;
    }
    t->last_queue_count = shadow_pass ? q.count : nullptr;
    unsigned long long *next1 = nullptr, *next2 = nullptr;
    rc = next_counter(t, cs, &next1);
    if (rc) return rc;
    if (shadow_pass) { rc = next_counter(t, cs, &next2); if (rc) return rc; }
    const bool fuse = t->fused_shade != 0; // vx_tracer_set_fused_shade: a frame in another GPU's memory is shaded inside the march
    rc = launch_frame(t, w, f, d_bgra, d_steps, st, q, next1, next2, opts->norm_mode == VX_NORM_HOST, shadow_pass && !t->primary_only, fuse, cs);
    if (rc) return rc;
    VX_CUDA(cudaGetLastError());
    return VX_OK;
}

// ---- vx_render_static: the shadowed frame of vx_render for callers that may not allocate -----------------------------------
// No counterpart in the reference; added (CUDA graph capture, serving loops and harnesses
// that forbid cudaMalloc).  Everything vx_render keeps in a tracer's cudaMalloc'd buffers — the three ray / queue counters and the
// queue of primary hits waiting for their sun ray (record, aux words, saved lane state: 72 bytes per ray id) — lives in ONE
// file-scope device array of fixed size; every GPU that loads the library holds it until exit.  The kernels are vx_render's own.
constexpr unsigned long long kStaticMaxIds = VX_STATIC_MAX_RAY_IDS; // ray ids per call (128 per 16x8 pixel tile: 640x360 = 230400)
constexpr size_t kStaticHeader = 256;                               // counters: frame queue count, primary ray counter, sun ray counter
constexpr size_t kStaticRecOff = kStaticHeader;
constexpr size_t kStaticAuxOff = kStaticRecOff + (size_t)kStaticMaxIds * sizeof(float4);
constexpr size_t kStaticLanesOff = kStaticAuxOff + (size_t)kStaticMaxIds * 2u * sizeof(uint32_t);
constexpr size_t kStaticBytes = kStaticLanesOff + (size_t)kStaticMaxIds * 3u * sizeof(uint4);
static __device__ __align__(256) unsigned char g_static_scratch[kStaticBytes];

namespace {
struct StaticDevice { unsigned char* base = nullptr; int sm_count = 0; bool ready = false; };
constexpr int kStaticMaxDevices = 64;
StaticDevice g_static_device[kStaticMaxDevices];
std::mutex g_static_mu;

// Once per device, with that device current: in host code `g_static_scratch` is only a stand-in for the symbol, the address the
// kernels need comes from cudaGetSymbolAddress.
int static_setup(int device, StaticDevice& out) {
    std::lock_guard<std::mutex> lock(g_static_mu);
    VX_REQUIRE(device >= 0 && device < kStaticMaxDevices, "vx_render_static: device index out of range");
    StaticDevice& d = g_static_device[device];
    if (!d.ready) {
        void* p = nullptr;
        VX_CUDA(cudaGetSymbolAddress(&p, g_static_scratch));
        VX_REQUIRE(p != nullptr, "vx_render_static: cudaGetSymbolAddress returned NULL");
        VX_CUDA(cudaDeviceGetAttribute(&d.sm_count, cudaDevAttrMultiProcessorCount, device));
        d.base = (unsigned char*)p;
        d.ready = true;
    }
    out = d;
    return VX_OK;
}
} // namespace

extern "C" int vx_render_static(const vx_world* world, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts, void* d_bgra,
                                int32_t* d_steps, uint32_t W, uint32_t H, void* stream) {
    VX_REQUIRE(world && cam && env && opts && d_bgra, "NULL argument");
    VX_REQUIRE(opts->norm_mode == VX_NORM_DEVICE || opts->norm_mode == VX_NORM_HOST, "render supports VX_NORM_DEVICE or VX_NORM_HOST");
    VX_REQUIRE(opts->d_stats == nullptr, "vx_render_static: the counting kernels (d_stats) need a tracer; use vx_render");
    VX_REQUIRE(cam->tan_half > 0.0f, "vx_render_static: pass cam->tan_half (vx_device_tanf, once at set-up); evaluating it here would allocate");
    DeviceGuard guard(world->device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    StaticDevice sd;
    int rc = static_setup(world->device, sd);
    if (rc) return rc;
    vx_tracer local;  // plain host struct (no device state): what fill_frame and persistent_blocks read
    local.world = const_cast<vx_world*>(world); local.device = world->device; local.sm_count = sd.sm_count;
    FrameParams f;
    rc = fill_frame(&local, cam, env, opts, W, H, f);
    if (rc) return rc;
    if (f.tw == 0 || f.th == 0) return VX_OK;
    const uint32_t bands = (f.th + 7u) / 8u;
    const uint32_t my_bands = bands > f.band_index ? (bands - f.band_index + f.band_count - 1u) / f.band_count : 0u;
    if (my_bands == 0) return VX_OK;
#if VX_TILE_SHAPE == 2
    f.mcols = (f.tw + 31u) / 32u * 2u;
#else
    f.mcols = (f.tw + 15u) / 16u;
#endif
    f.ids = (unsigned long long)f.mcols * my_bands * 128ull;
    if (f.ids > kStaticMaxIds) {
        char msg[256];
        snprintf(msg, sizeof msg, "vx_render_static: this frame needs %llu ray ids, the static scratch holds VX_STATIC_MAX_RAY_IDS = %llu "
                 "(e.g. 640x360); render it in tiles or bands, or use vx_render", f.ids, kStaticMaxIds);
        set_error(msg);
        return VX_ERR_INVALID;
    }
    f.tile_trace = nullptr;
    const WorldView w = world->view();
    cudaStream_t cs = (cudaStream_t)stream;
    // the array keeps the last call's contents: the three counters must start at zero, the queue entries are written before read
    VX_CUDA(cudaMemsetAsync(sd.base, 0, kStaticHeader, cs));
    unsigned long long* counters = (unsigned long long*)sd.base;
    ShadowQueue q{nullptr, nullptr, nullptr, nullptr};
    const bool shadow_pass = opts->shadows && !opts->debug_view;
    if (shadow_pass) {
        q.rec = (float4*)(sd.base + kStaticRecOff); q.aux = (uint32_t*)(sd.base + kStaticAuxOff); q.lanes = (uint4*)(sd.base + kStaticLanesOff);
        q.count = counters;
    }
    unsigned long long *next1 = counters + 8, *next2 = counters + 16;
    rc = launch_frame(&local, w, f, d_bgra, d_steps, nullptr, q, next1, next2, opts->norm_mode == VX_NORM_HOST, shadow_pass, false, cs);
    if (rc) return rc;
    VX_CUDA(cudaGetLastError());
    return VX_OK;
}

extern "C" int vx_bounce_rays(vx_tracer* t, const vx_camera* cam, const vx_render_opts* opts, uint32_t W, uint32_t H, float* d_origins,
                              float* d_dirs, uint32_t* d_pixels, unsigned long long* d_count, void* stream) {
    VX_REQUIRE(t && cam && opts && d_origins && d_dirs && d_pixels && d_count, "NULL argument");
    VX_REQUIRE(opts->norm_mode == VX_NORM_DEVICE || opts->norm_mode == VX_NORM_HOST, "bounce rays support VX_NORM_DEVICE or VX_NORM_HOST");
    DeviceGuard guard(t->device);
    cudaStream_t cs = (cudaStream_t)stream;
    // the primary pass of a shadowed frame, into the tracer's own frame buffer; its hit queue is the input
    const size_t bytes = (size_t)W * H * 4;
    if (t->frame_bytes < bytes) {
        VX_CUDA(cudaStreamSynchronize(cs));
        cudaFree(t->d_frame); t->d_frame = nullptr; t->frame_bytes = 0;
        VX_CUDA(cudaMalloc(&t->d_frame, bytes));
        t->frame_bytes = bytes; t->frame_seeded = false;
    }
    vx_render_opts o = *opts;
    o.shadows = 1; o.debug_view = 0; o.d_stats = nullptr;
    vx_env env = {{1.0f, 1.0f, 1.0f}, {0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}};
    t->primary_only = true;
    int rc = vx_render(t, cam, &env, &o, t->d_frame, nullptr, W, H, stream);
    t->primary_only = false;
    if (rc) return rc;
    FrameParams f;
    rc = fill_frame(t, cam, &env, &o, W, H, f);
    if (rc) return rc;
    if (!t->last_queue_count) { VX_CUDA(cudaMemsetAsync(d_count, 0, sizeof(unsigned long long), cs)); return VX_OK; }
    ShadowQueue q{(float4*)t->d_queue_rec, (uint32_t*)t->d_queue_aux, (uint4*)t->d_queue_lanes, t->last_queue_count};
    const unsigned dense_blocks = (unsigned)(t->sm_count > 0 ? t->sm_count : 148) * 8u;
    if (o.norm_mode == VX_NORM_HOST) bounce_setup<1><<<dense_blocks, 256, 0, cs>>>(f, q, d_origins, d_dirs, d_pixels);
    else bounce_setup<0><<<dense_blocks, 256, 0, cs>>>(f, q, d_origins, d_dirs, d_pixels);
    VX_CUDA(cudaGetLastError());
    VX_CUDA(cudaMemcpyAsync(d_count, t->last_queue_count, sizeof(unsigned long long), cudaMemcpyDeviceToDevice, cs));
    return VX_OK;
}

extern "C" int vx_tracer_set_fused_shade(vx_tracer* t, int mode) {
    VX_REQUIRE(t != nullptr, "tracer is NULL");
    VX_REQUIRE(mode == 0 || mode == 1, "mode is 0 or 1");
    t->fused_shade = mode;
    return VX_OK;
}

extern "C" int vx_tracer_set_timing(vx_tracer* t, int enable) {
    VX_REQUIRE(t != nullptr, "NULL argument");
    DeviceGuard guard(t->device);
    if (enable && !t->ev[0]) for (int i = 0; i < 5; ++i) VX_CUDA(cudaEventCreate(&t->ev[i]));
    t->timing = enable != 0;
    t->timed_kernels = 0;
    return VX_OK;
}

extern "C" int vx_tracer_frame_times(vx_tracer* t, float ms[4]) {
    VX_REQUIRE(t && ms, "NULL argument");
    VX_REQUIRE(t->timing && t->timed_kernels > 0, "no timed frame: call vx_tracer_set_timing(t, 1) and render first");
    DeviceGuard guard(t->device);
    for (int i = 0; i < 4; ++i) ms[i] = 0.0f;
    VX_CUDA(cudaEventSynchronize(t->ev[t->timed_kernels]));
    for (int i = 0; i < t->timed_kernels; ++i) VX_CUDA(cudaEventElapsedTime(&ms[i], t->ev[i], t->ev[i + 1]));
    return VX_OK;
}

extern "C" int vx_render_host(vx_tracer* t, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts, void* h_bgra, uint32_t W,
                              uint32_t H) {
    VX_REQUIRE(t && h_bgra, "NULL argument");
    DeviceGuard guard(t->device);
    const size_t bytes = (size_t)W * H * 4;
    if (t->frame_bytes < bytes) {
        cudaFree(t->d_frame); t->d_frame = nullptr; t->frame_bytes = 0;
        VX_CUDA(cudaMalloc(&t->d_frame, bytes));
        t->frame_bytes = bytes; t->frame_seeded = false;
    }
    if (!t->frame_seeded) {
        t->frame_seeded = true;
        // first use: seed the device frame with the caller's buffer so that unwritten pixels
        // (checkerboard, debug bottom-left) keep their previous value, as in the reference
        VX_CUDA(cudaMemcpyAsync(t->d_frame, h_bgra, bytes, cudaMemcpyHostToDevice, t->stream));
    }
    // Large plain frames are rendered in horizontal slices so that the copy of one slice to the host overlaps the render of
    // the next (second stream, one event per slice).  Two uneven slices measured best at 4K: the first 80 % of the rows, then
    // the rest, so that only a fifth of the frame is still to be copied when the last kernel ends (ms per frame: one pass
    // 3.44, 50/50 3.33, 70/30 3.18, 80/20 3.10, 90/10 3.12; three even slices 3.47 — more slices lose more to launch tails
    // than they hide; letting the kernels store pixels straight into pinned host memory over PCIe is far slower).  VX_HOST_SLICES=1..4 overrides.  Checkerboard frames (rows of neighbouring threads interleave) and
    // callers that already partition the frame take the single-pass path.
    static const uint32_t kSlices = [] { const char* e = getenv("VX_HOST_SLICES"); int v = e ? atoi(e) : 2; return (uint32_t)(v < 1 ? 1 : (v > 4 ? 4 : v)); }();
    const bool sliced = opts && !opts->checkerboard && opts->tile_w == 0 && opts->tile_h == 0 && opts->band_count == 0 && kSlices > 1 && H >= 64u * kSlices;
    if (!sliced) {
        int rc = vx_render(t, cam, env, opts, t->d_frame, nullptr, W, H, t->stream);
        if (rc) return rc;
        VX_CUDA(cudaMemcpyAsync(h_bgra, t->d_frame, bytes, cudaMemcpyDeviceToHost, t->stream));
        VX_CUDA(cudaStreamSynchronize(t->stream));
        return VX_OK;
    }
    if (!t->copy_stream) {
        VX_CUDA(cudaStreamCreateWithFlags(&t->copy_stream, cudaStreamNonBlocking));
        for (uint32_t k = 0; k < kSlices; ++k) VX_CUDA(cudaEventCreateWithFlags(&t->slice_ev[k], cudaEventDisableTiming));
    }
    static const float kSplit = [] { const char* e = getenv("VX_HOST_SPLIT"); float v = e ? (float)atof(e) : 0.8f; return v < 0.1f ? 0.1f : (v > 0.97f ? 0.97f : v); }();
    const uint32_t even_rows = ((H + kSlices - 1) / kSlices + 7u) / 8u * 8u;
    for (uint32_t k = 0, y0 = 0, rows = 0; y0 < H; ++k, y0 += rows) {
        // two slices: the first takes kSplit of the rows, so that less of the frame is left to copy after the last kernel
        rows = kSlices == 2 ? (k == 0 ? ((uint32_t)((float)H * kSplit) + 7u) / 8u * 8u : H - y0) : even_rows;
        vx_render_opts o = *opts;
        o.tile_x0 = 0; o.tile_y0 = y0; o.tile_w = (W + 31u) / 32u * 32u; o.tile_h = rows;
        int rc = vx_render(t, cam, env, &o, t->d_frame, nullptr, W, H, t->stream);
        if (rc) return rc;
        const uint32_t h = y0 + rows <= H ? rows : H - y0;
        VX_CUDA(cudaEventRecord(t->slice_ev[k], t->stream));
        VX_CUDA(cudaStreamWaitEvent(t->copy_stream, t->slice_ev[k], 0));
        VX_CUDA(cudaMemcpyAsync((uint8_t*)h_bgra + (size_t)y0 * W * 4, (uint8_t*)t->d_frame + (size_t)y0 * W * 4, (size_t)h * W * 4,
                                cudaMemcpyDeviceToHost, t->copy_stream));
    }
    VX_CUDA(cudaStreamSynchronize(t->copy_stream));
    VX_CUDA(cudaStreamSynchronize(t->stream));
    return VX_OK;
}

// ---- diagnostics: streaming-read bandwidth (L2-resident when `bytes` is well under the 126 MB L2) ----
namespace vx {
__global__ void __launch_bounds__(512) stream_read(const uint4* __restrict__ buf, size_t n16, int reps, unsigned* __restrict__ sink) {
    unsigned acc = 0;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    for (int r = 0; r < reps; ++r)
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += stride) {
            uint4 v;
            asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(buf + i));
            acc ^= v.x ^ v.y ^ v.z ^ v.w;
        }
    if (acc == 0x9E3779B9u) *sink = acc; // never true for a zero buffer; keeps the loads alive
}
} // namespace vx

extern "C" int vx_measure_read_bandwidth(int device, size_t bytes, int reps, float* gb_per_s) {
    VX_REQUIRE(gb_per_s && bytes >= 4096 && reps > 0, "bad argument");
    DeviceGuard guard(device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    void* buf = nullptr; unsigned* sink = nullptr;
    VX_CUDA(cudaMalloc(&buf, bytes));
    VX_CUDA(cudaMalloc(&sink, 4));
    VX_CUDA(cudaMemset(buf, 0, bytes));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    cudaEvent_t e0, e1;
    VX_CUDA(cudaEventCreate(&e0)); VX_CUDA(cudaEventCreate(&e1));
    float best = 0.0f;
    for (int trial = 0; trial < 5; ++trial) {
        stream_read<<<sms * 4, 512>>>((const uint4*)buf, bytes / 16, trial == 0 ? 1 : reps, sink); // trial 0 warms the cache
        if (trial == 0) continue;
        VX_CUDA(cudaEventRecord(e0));
        stream_read<<<sms * 4, 512>>>((const uint4*)buf, bytes / 16, reps, sink);
        VX_CUDA(cudaEventRecord(e1));
        VX_CUDA(cudaEventSynchronize(e1));
        float ms = 0.0f;
        VX_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const float g = (float)((double)(bytes / 16 * 16) * reps / (ms * 1e-3) / 1e9);
        if (g > best) best = g;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(buf); cudaFree(sink);
    *gb_per_s = best;
    return VX_OK;
}

#ifdef VX_FINE_STATS
// diagnostics build only: [0] quick marches started, [1] start cell out of range, [2] two far faces, [3] far face not first,
// [4] crossing test inconclusive, [5] starts on a far face; reading resets the counters
extern "C" int vx_debug_fine_stats(unsigned long long out[8], int device) {
    DeviceGuard guard(device);
    VX_CUDA(cudaDeviceSynchronize());
    VX_CUDA(cudaMemcpyFromSymbol(out, vx::g_fine_stats, 64));
    unsigned long long zero[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    VX_CUDA(cudaMemcpyToSymbol(vx::g_fine_stats, zero, 64));
    return VX_OK;
}
#endif

// diagnostics (VX_TILE_TRACE=1): for every 8x4 tile of the last vx_render, the SM cycles it kept its warp and the low word of
// %globaltimer (ns) at which the warp drew it; out arrays hold `capacity` entries, *n = tiles of that launch
// Diagnostics (not in vxrt.h): the march's two exactness-critical shortcuts evaluated on the device for host arrays, so a test
// can put them beside the operations they stand for — tmax_div(a, d) against a / d, next_toward(x, down) against
// nextafterf(x, -+inf).  out: 6 planes of n floats — tmax_div(a, d), __fdiv_rn(a, d), next_toward(a, up), next_toward(a, down),
// nextafterf(a, +inf), nextafterf(a, -inf).
__global__ void chain_math_kernel(const float* __restrict__ a, const float* __restrict__ d, size_t n, float* __restrict__ out) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    out[i] = vx::tmax_div(a[i], d[i]);
    out[n + i] = __fdiv_rn(a[i], d[i]);
    out[2 * n + i] = vx::next_toward(a[i], false);
    out[3 * n + i] = vx::next_toward(a[i], true);
    out[4 * n + i] = nextafterf(a[i], CUDART_INF_F);
    out[5 * n + i] = nextafterf(a[i], -CUDART_INF_F);
}
extern "C" int vx_debug_chain_math(const float* a, const float* d, size_t n, float* out, int device) {
    VX_REQUIRE(a && d && out, "NULL argument");
    if (!n) return VX_OK;
    DeviceGuard guard(device);
    float* buf = nullptr;
    VX_CUDA(cudaMalloc(&buf, n * 32));
    cudaError_t e = cudaMemcpy(buf, a, n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(buf + n, d, n * 4, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) {
        chain_math_kernel<<<(unsigned)((n + 255) / 256), 256>>>(buf, buf + n, n, buf + 2 * n);
        e = cudaMemcpy(out, buf + 2 * n, n * 24, cudaMemcpyDeviceToHost);
    }
    cudaFree(buf);
    VX_CUDA(e);
    return VX_OK;
}

extern "C" int vx_debug_tile_trace(vx_tracer* t, uint32_t* out_cycles, uint32_t* out_start_ns, size_t capacity, size_t* n) {
    VX_REQUIRE(t && n, "NULL argument");
    *n = 0;
    if (!t->d_tile_trace) return VX_OK;
    DeviceGuard guard(t->device);
    VX_CUDA(cudaDeviceSynchronize());
    const size_t m = t->tile_trace_tiles < capacity ? t->tile_trace_tiles : capacity;
    if (out_cycles) VX_CUDA(cudaMemcpy(out_cycles, t->d_tile_trace, m * 4, cudaMemcpyDeviceToHost));
    if (out_start_ns) VX_CUDA(cudaMemcpy(out_start_ns, t->d_tile_trace + t->tile_trace_tiles, m * 4, cudaMemcpyDeviceToHost));
    *n = m;
    return VX_OK;
}
