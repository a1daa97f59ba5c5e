// vx_march.cuh — the two-level march as a per-lane state machine driven by persistent warps.
//
// Same arithmetic as vx_device.cuh's primitives (VolumeRaytracer.cu:124-174, :176-352, :354-525 under IEEE
// binary32 without contraction); what is new is the control structure:
//   * a ray is a `Lane` record that moves between states BEGIN -> FAST/SLOW -> CAND -> (BEGIN | DONE);
//   * a warp runs ONE loop whose body has one section per state, so lanes that are in the same state execute
//     together no matter which ray, which coarse traversal or which phase (primary / shadow) they belong to;
//   * finished lanes are refilled from a global ray counter (persistent warps), and a primary hit turns into its
//     shadow ray in place;
//   * the step counter of a coarse traversal is not carried: every step moves one cell along one axis in a fixed
//     direction, so steps = L1 distance between the current cell and the start cell.
#pragma once
#include "vx_device.cuh"

namespace vx {

constexpr uint32_t kNoCell = 0xFFFFFFFFu;
constexpr uint32_t kFull = 0xFFFFFFFFu;

enum : uint32_t { ST_IDLE = 0, ST_BEGIN = 1, ST_FAST = 2, ST_SLOW = 3, ST_CAND = 4, ST_DONE = 5 };
constexpr uint32_t F_ST = 7u;          // bits 0-2: state
constexpr uint32_t F_SLOW = 1u << 3;   // traversal runs in the generic (bounds-checked, reference-ordered) mode
constexpr uint32_t F_EDGE = 1u << 4;   // IsOnEdge at the start of this traversal (VolumeRaytracer.cu:216-232)
constexpr uint32_t F_CAPOK = 1u << 5;  // no tMax is NaN: the fast loop's step is the reference's comparison chain
constexpr uint32_t F_HIT = 1u << 6;
constexpr uint32_t F_SHADOW = 1u << 7; // render: this lane marches the shadow ray of its pixel
constexpr int F_SN_SHIFT = 8;          // bits 8-10: normal code of the grid-entry clip (start_normal)
constexpr uint32_t F_NEAR = 1u << 11;  // the kMaxSteps cap can be reached before the border: the fast loop checks per chunk
constexpr uint32_t F_DNZ = 1u << 12;   // every direction component is finite and not zero: the tMax divisions need no guard
constexpr uint32_t F_DPOS = 1u << 13;  // ... and positive (the sun rays of the reference's light): then face - s = trunc(s) + 1 - s is never zero either

// Round-2 trims of the per-brick chain (slab test -> brick march -> re-seed -> traversal start); each can be switched off
// (-DVX_OPT_x=0) to measure it.  None changes a result bit.
#ifndef VX_OPT_TDIV
#define VX_OPT_TDIV 1  // tMax divisions: no d != 0 guard for F_DNZ rays, zero numerators answered without div.rn's slow path
#endif
#ifndef VX_OPT_NEXT
#define VX_OPT_NEXT 1  // nextafterf(x, +-inf) of the re-seeding nudge as an integer step on the bits
#endif
#ifndef VX_OPT_CSTEP
#define VX_OPT_CSTEP 1 // the step after a slab miss of a fast-mode traversal: the fast loop's predicated step
#endif
#ifndef VX_OPT_FMA
#define VX_OPT_FMA 1   // tight-box corners as one exact FMA each
#endif
#ifndef VX_OPT_DPOS
#define VX_OPT_DPOS 1  // F_DPOS rays: plain divisions
#endif

// Axis-aligned unit normals as a 3-bit code: 0 = (0,0,0); 1 + 2*axis + (negative ? 1 : 0).
__device__ __forceinline__ V3 normal_from_code(uint32_t c) {
    const float v = (c & 1u) ? 1.0f : -1.0f; // codes 1,3,5 are positive
    return mk((c == 1u || c == 2u) ? v : 0.0f, (c == 3u || c == 4u) ? v : 0.0f, (c == 5u || c == 6u) ? v : 0.0f);
}

struct Lane {
    V3 d;                // normalised direction
    float ivx, ivy, ivz; // slab inverse 1/(d == 0 ? eps : d)
    V3 s;                // start of the current coarse traversal (brick units)
    float tx, ty, tz;    // tMax of the current coarse traversal
    uint32_t idx;        // bordered linear index of the current coarse cell
    uint32_t bidx;       // ST_CAND: bordered index of the occupied cell to test (clamped cell in the generic mode)
    uint32_t pc;         // bordered index of the previous candidate brick (VolumeRaytracer.cu:402-407) or kNoCell
    uint32_t flags;
    int total;           // total_steps
    // batch planes / work counters (dead code in the render kernel unless it counts)
    int csteps, fsteps;
    uint32_t fcell;      // fine cell of the hit, packed 10 bits per axis
    Work work;
};

__device__ __forceinline__ uint32_t lane_state(const Lane& L) { return L.flags & F_ST; }
__device__ __forceinline__ void lane_set_state(Lane& L, uint32_t st) { L.flags = (L.flags & ~F_ST) | st; }

__device__ __forceinline__ float lane_dt(float d, float iv) { return d != 0.0f ? fabsf(iv) : CUDART_INF_F; } // :199-201
__device__ __forceinline__ Ray lane_ray(const Lane& L) {
    Ray r;
    r.dnz = (L.flags & F_DNZ) != 0u; r.dpos = (L.flags & F_DPOS) != 0u;
    r.d = L.d; r.ivx = L.ivx; r.ivy = L.ivy; r.ivz = L.ivz;
    r.dtx = lane_dt(L.d.x, L.ivx); r.dty = lane_dt(L.d.y, L.ivy); r.dtz = lane_dt(L.d.z, L.ivz);
    r.sx = L.d.x > 0.0f ? 1 : -1; r.sy = L.d.y > 0.0f ? 1 : -1; r.sz = L.d.z > 0.0f ? 1 : -1;
    return r;
}

// bordered index <-> padded coordinates (px = x + 1, ...); divisions by the two pitches use host-made magic numbers
__device__ __forceinline__ void cell_of(const WorldView& w, uint32_t idx, int& x, int& y, int& z) {
    const uint32_t pz = __umulhi(idx, w.magic_xy) >> w.shift_xy;
    const uint32_t rem = idx - pz * (uint32_t)w.pnxy;
    const uint32_t py = __umulhi(rem, w.magic_x) >> w.shift_x;
    x = (int)(rem - py * (uint32_t)w.pnx) - 1; y = (int)py - 1; z = (int)pz - 1;
}
__device__ __forceinline__ uint32_t index_of(const WorldView& w, int x, int y, int z) {
    return (uint32_t)((z + 1) * w.pnxy + (y + 1) * w.pnx + (x + 1));
}
__device__ __forceinline__ bool interior(const WorldView& w, int x, int y, int z) {
    return (unsigned)x < (unsigned)w.cx && (unsigned)y < (unsigned)w.cy && (unsigned)z < (unsigned)w.cz;
}
__device__ __forceinline__ int steps_from_start(const Lane& L, int x, int y, int z) {
    return abs(x - (int)L.s.x) + abs(y - (int)L.s.y) + abs(z - (int)L.s.z);
}

// p / factor (start = origin / factor, VolumeRaytracer.cu:361, and the re-seeding of :438).  For a power-of-two factor
// x * (1 / factor) is the same float as x / factor — both are the correctly rounded value of the same real number.
__device__ __forceinline__ V3 to_brick_units(const WorldView& w, V3 p) {
    const float inv_f = w.inv_factor;
    if (inv_f != 0.0f) return mk(fmul(p.x, inv_f), fmul(p.y, inv_f), fmul(p.z, inv_f));
    const float ff = (float)w.factor;
    return mk(fdiv(p.x, ff), fdiv(p.y, ff), fdiv(p.z, ff));
}

// tMax = (face - s) / d (VolumeRaytracer.cu:203-205) for a direction component that is finite and not zero (F_DNZ).
// A zero numerator — the start lies exactly on the face it is measured to: every brick entered through an upper face of
// its tight box by a ray that descends along that axis, every camera on an integer brick coordinate — gives (+-0) / d, a
// zero with the XOR of the signs.  div.rn's in-line path does not take zero operands (FCHK sends them through a
// 40-instruction routine, 2.7 % of primary_rays' instructions at the benchmark camera), so that case is answered here and
// the division runs on a stand-in numerator.
__device__ __forceinline__ float tmax_div(float a, float d) {
    const bool z = a == 0.0f;
    const float q = fdiv(z ? 1.0f : a, d);
    return z ? __uint_as_float((__float_as_uint(a) ^ __float_as_uint(d)) & 0x80000000u) : q;
}
__device__ __forceinline__ bool finite_nonzero(float v) { return (__float_as_uint(v) & 0x7FFFFFFFu) - 1u < 0x7F7FFFFFu; }

// nextafterf(x, down ? -inf : +inf) (the nudge of VolumeRaytracer.cu:452-460).  For a finite x that is neither zero nor
// the largest float it is one step on the bit pattern, away from zero or towards it; everything else goes to the library.
__device__ __forceinline__ float next_toward(float x, bool down) {
#if VX_OPT_NEXT
    const uint32_t b = __float_as_uint(x);
    if ((b & 0x7FFFFFFFu) - 1u < 0x7F7FFFFEu) return __uint_as_float(down == (bool)(b >> 31) ? b + 1u : b - 1u);
#endif
    return nextafterf(x, down ? -CUDART_INF_F : CUDART_INF_F);
}

// Ray set-up shared by primary, shadow and batch rays: slab inverses, grid-entry clip (VolumeRaytracer.cu:361-381).
// `s` is origin / factor, `d` the normalised direction.
__device__ __forceinline__ void lane_setup(const WorldView& w, Lane& L, V3 s, V3 d, uint32_t keep_flags) {
    L.d = d;
    L.ivx = fdiv(1.0f, d.x == 0.0f ? kFltEps : d.x);
    L.ivy = fdiv(1.0f, d.y == 0.0f ? kFltEps : d.y);
    L.ivz = fdiv(1.0f, d.z == 0.0f ? kFltEps : d.z);
    uint32_t sn = 0;
    if (!(s.x >= 0.0f && s.y >= 0.0f && s.z >= 0.0f && s.x < (float)w.cx && s.y < (float)w.cy && s.z < (float)w.cz)) {
        const float e = (float)kEpsDDA;
        const Ray r = lane_ray(L);
        const Slab sl = ray_aabb(s, r, mk(e, e, e),
                                 mk((float)((double)w.cx - kEpsDDA), (float)((double)w.cy - kEpsDDA), (float)((double)w.cz - kEpsDDA)));
        if (sl.hit) { s = sl.point; sn = sl.ncode; }
    }
    L.s = s;
    L.pc = kNoCell;
    L.total = 0; L.csteps = 0; L.fsteps = 0; L.fcell = 0;
    L.work.C = L.work.K = L.work.E = 0;
    L.flags = keep_flags | (sn << F_SN_SHIFT) | ST_BEGIN;
    if (VX_OPT_TDIV && finite_nonzero(d.x) && finite_nonzero(d.y) && finite_nonzero(d.z))
        L.flags |= F_DNZ | ((VX_OPT_DPOS && d.x > 0.0f && d.y > 0.0f && d.z > 0.0f) ? F_DPOS : 0u);
}

// Start of one coarse DDARayTraversal (VolumeRaytracer.cu:186-232) from L.s; also the `while (total_steps <
// maxSteps)` test of Raytrace (:383).
// SUN: the caller's rays share one direction (sun rays), so the F_DPOS test is warp-uniform and worth a code path of its own;
// kernels with mixed directions do not carry it (compiled into primary_rays it costs 5 % without ever being taken).
template <bool SUN>
__device__ __forceinline__ void lane_begin(const WorldView& w, Lane& L, int max_steps) {
    L.flags &= ~(F_SLOW | F_EDGE | F_CAPOK | F_NEAR);
    if (!(L.total < max_steps)) { lane_set_state(L, ST_DONE); return; }
    const int x0 = (int)L.s.x, y0 = (int)L.s.y, z0 = (int)L.s.z; // cvt.rzi: truncation, saturating, NaN -> 0
    const int ox = L.d.x > 0.0f, oy = L.d.y > 0.0f, oz = L.d.z > 0.0f;
    if (VX_OPT_DPOS && SUN && (L.flags & F_DPOS)) {
        L.tx = fdiv(fsub((float)(x0 + 1), L.s.x), L.d.x);
        L.ty = fdiv(fsub((float)(y0 + 1), L.s.y), L.d.y);
        L.tz = fdiv(fsub((float)(z0 + 1), L.s.z), L.d.z);
    } else if (L.flags & F_DNZ) {
        L.tx = tmax_div(fsub((float)(x0 + ox), L.s.x), L.d.x);
        L.ty = tmax_div(fsub((float)(y0 + oy), L.s.y), L.d.y);
        L.tz = tmax_div(fsub((float)(z0 + oz), L.s.z), L.d.z);
    } else {
        L.tx = L.d.x != 0.0f ? fdiv(fsub((float)(x0 + ox), L.s.x), L.d.x) : CUDART_INF_F;
        L.ty = L.d.y != 0.0f ? fdiv(fsub((float)(y0 + oy), L.s.y), L.d.y) : CUDART_INF_F;
        L.tz = L.d.z != 0.0f ? fdiv(fsub((float)(z0 + oz), L.s.z), L.d.z) : CUDART_INF_F;
    }
    // cells in [-1, dim] have a bordered index; anything else fails the first bounds test with zero steps (:283-287)
    if (!((unsigned)(x0 + 1) <= (unsigned)(w.cx + 1) && (unsigned)(y0 + 1) <= (unsigned)(w.cy + 1) &&
          (unsigned)(z0 + 1) <= (unsigned)(w.cz + 1))) { lane_set_state(L, ST_DONE); return; }
    L.idx = index_of(w, x0, y0, z0);
    const bool on_edge = x0 == w.cx || y0 == w.cy || z0 == w.cz;
    const bool no_nan = L.tx == L.tx && L.ty == L.ty && L.tz == L.tz;
    // examinations needed to reach the border along every axis: the fast loop carries no step counter, so where the
    // kMaxSteps cap (:236) could come before a border cell (which ends the traversal) the loop re-derives the count
    // from the cell index before every chunk of kFastChunk steps (F_NEAR) and hands the last stretch to lane_slow
    const int reach = (ox ? w.cx - x0 : x0 + 1) + (oy ? w.cy - y0 : y0 + 1) + (oz ? w.cz - z0 : z0 + 1);
    const bool capok = no_nan;
    const bool fast = capok && interior(w, x0, y0, z0);
    L.flags |= (on_edge ? F_EDGE : 0u) | (capok ? F_CAPOK : 0u) | (fast ? 0u : F_SLOW) | (reach < kMaxSteps ? 0u : F_NEAR);
    lane_set_state(L, fast ? ST_FAST : ST_SLOW);
}

// One cell along the smallest tMax in the reference's comparison order (ties: y over x, z over both — :293-322).
// Returns the cell after the step through (nx,ny,nz) deltas applied to idx.
__device__ __forceinline__ void lane_step(const WorldView& w, Lane& L) {
    const bool mx = L.tx < L.ty && L.tx < L.tz;
    const bool my = !mx && L.ty <= L.tx && L.ty < L.tz;
    const bool mz = !mx && !my;
    const float t = mx ? L.tx : (my ? L.ty : L.tz);
    const float dt = mx ? lane_dt(L.d.x, L.ivx) : (my ? lane_dt(L.d.y, L.ivy) : lane_dt(L.d.z, L.ivz));
    const float tn = fadd(t, dt);
    L.tx = mx ? tn : L.tx; L.ty = my ? tn : L.ty; L.tz = mz ? tn : L.tz;
    const int dix = L.d.x > 0.0f ? 1 : -1, diy = L.d.y > 0.0f ? w.pnx : -w.pnx, diz = L.d.z > 0.0f ? w.pnxy : -w.pnxy;
    L.idx += (uint32_t)(mx ? dix : (my ? diy : diz));
}

// One fast-loop step.  With no NaN among the tMax values (F_CAPOK) the reference's comparison chain (:293-322) picks the
// axis of the minimum with ties z over y over x, and the chosen axis' tMax IS that minimum, so "tMax[axis] += tDelta[axis]"
// under the axis predicate is the reference's add.  Written as predicated adds (FMA pipe) because the select-based form
// the compiler prefers is bound by the half-rate ALU pipe (profiles/).
__device__ __forceinline__ void fast_step(float& tx, float& ty, float& tz, uint32_t& idx, float dtx, float dty, float dtz, int dix, int diy,
                                          int diz) {
    asm("{\n\t.reg .pred pz, py, px;\n\t.reg .f32 m;\n\t"
        "min.f32 m, %0, %1;\n\t"
        "setp.le.f32 pz, %2, m;\n\t"
        "setp.le.and.f32 py|px, %1, %0, !pz;\n\t" // py = (ty <= tx) & !pz, px = !(ty <= tx) & !pz
        "@pz add.rn.f32 %2, %2, %6;\n\t"
        "@py add.rn.f32 %1, %1, %5;\n\t"
        "@px add.rn.f32 %0, %0, %4;\n\t"
        "@pz add.s32 %3, %3, %9;\n\t"
        "@py add.s32 %3, %3, %8;\n\t"
        "@px add.s32 %3, %3, %7;\n\t}"
        : "+f"(tx), "+f"(ty), "+f"(tz), "+r"(idx)
        : "f"(dtx), "f"(dty), "f"(dtz), "r"(dix), "r"(diy), "r"(diz));
}

#ifndef VX_REFILL_MIN
#define VX_REFILL_MIN 8
#endif
#ifndef VX_FAST_CHUNK
#define VX_FAST_CHUNK 64
#endif
// Fast-loop scheduling, per ray source (Policy::kFastQuad, kLeaveNum / kLeaveDen): steps between looks at how many lanes are
// still stepping, and the share of the lanes that entered below which the rest stop too.  Measured per 4K frame (profiles/):
// primaries 4 steps, 1/4: 1.41 ms; 8 steps: 1.37 (1/4 and 1/2 alike); sun rays 4 steps, 1/4: 1.00 ms; 4 steps, 1/2: 0.95-0.96;
// incoherent bounce rays (batch kernel) 4 steps, 1/4: 2.30 ms; 8 steps: 2.25; 1/2: 2.46.
#ifndef VX_PRIMARY_QUAD
#define VX_PRIMARY_QUAD 4 // round 2, after the chain trims: 4 steps 1.280 / 0.594 / 0.527 ms (overview / grazing / diagonal camera), 8 steps
                          // 1.283 / 0.610 / 0.545, 2 steps 1.321 / 0.593 / 0.525 (profiles/r02_sweep_fast_loop_schedule_after_trims.log)
#endif
#ifndef VX_PRIMARY_LEAVE_NUM
#define VX_PRIMARY_LEAVE_NUM 1
#define VX_PRIMARY_LEAVE_DEN 2
#endif
#ifndef VX_SHADOW_QUAD
#define VX_SHADOW_QUAD 8 // round 2, after the brick march got cheaper: 8 steps 0.872 ms, 4 steps 0.888 (profiles/)
#endif
#ifndef VX_SHADOW_LEAVE_NUM
#define VX_SHADOW_LEAVE_NUM 1
#define VX_SHADOW_LEAVE_DEN 2
#endif
#ifndef VX_BATCH_QUAD
#define VX_BATCH_QUAD 8
#endif
#ifndef VX_BATCH_LEAVE_NUM
#define VX_BATCH_LEAVE_NUM 1
#define VX_BATCH_LEAVE_DEN 4
#endif
constexpr int kChunk = 32;                  // ray ids a warp draws at a time (render: one 8x4 pixel tile)
constexpr int kFastChunk = VX_FAST_CHUNK;   // coarse steps a lane may take before the warp looks at the other states again

// Generic mode: one bounds-checked examination + step per call (the ray starts outside the grid, sits on the padded
// far face, carries NaNs, or could reach the step cap).  Rare; exactness over speed.
template <bool COUNT>
__device__ __forceinline__ void lane_slow(const WorldView& w, Lane& L) {
    int x, y, z;
    cell_of(w, L.idx, x, y, z);
    const int it = steps_from_start(L, x, y, z);
    if ((L.flags & F_CAPOK) && interior(w, x, y, z) && (!(L.flags & F_NEAR) || it + kFastChunk <= kMaxSteps)) {
        L.flags &= ~F_SLOW; lane_set_state(L, ST_FAST); return;
    }
    const bool edge = L.flags & F_EDGE;
    const int limx = w.cx + ((edge && L.d.x < 0.0f) ? 1 : 0), limy = w.cy + ((edge && L.d.y < 0.0f) ? 1 : 0),
              limz = w.cz + ((edge && L.d.z < 0.0f) ? 1 : 0);
    if (it >= kMaxSteps || !((unsigned)x < (unsigned)limx && (unsigned)y < (unsigned)limy && (unsigned)z < (unsigned)limz)) {
        L.total += it; L.csteps += it; // cap reached or out of bounds: no hit (:236, :283-287)
        lane_set_state(L, ST_DONE);
        return;
    }
    const uint32_t b = index_of(w, min(x, w.cx - 1), min(y, w.cy - 1), min(z, w.cz - 1));
    const bool occ = (__ldg(w.ptop + (b >> 5)) >> (b & 31u)) & 1u;
    if (COUNT) { L.work.C++; L.work.K += occ; }
    if (occ) { L.bidx = b; lane_set_state(L, ST_CAND); return; }
    lane_step(w, L);
}

// Occupancy bit of bordered cell idx; the AND and the compare are one LOP3 with a predicate result.
__device__ __forceinline__ bool cell_bit(const uint32_t* __restrict__ bits, uint32_t idx) {
    const uint32_t word = __ldg(bits + (idx >> 5));
    uint32_t p;
    asm("{\n\t.reg .pred q;\n\t.reg .b32 t, u;\n\t"
        "shf.r.wrap.b32 t, %1, 0, %2;\n\t"
        "lop3.or.b32 u|q, t, 1, 0, 0xC0, 0;\n\t" // u = t & 1; q = (u != 0)
        "selp.u32 %0, 1, 0, q;\n\t}" : "=r"(p) : "r"(word), "r"(idx));
    return p != 0u;
}

// Fine level: DDARayTraversal inside one brick (cells = voxels), crossing points tested against [0,factor]^3 before a
// step is counted (VolumeRaytracer.cu:325-341).  FINE8: factor == 8, mask bit = x + 8y + 64z.
struct FineOut {
    bool hit, oob;
    int steps;
    uint32_t cell;   // packed 10 bits per axis
    uint32_t ncode;  // normal of the last counted step
    V3 point;
};
template <bool FINE8>
__device__ __forceinline__ void fine_march(const uint32_t* __restrict__ mask, int f, V3 s, const Ray& r, FineOut& o) {
    Dda q = dda_begin(s, r, f, f, f);
    o.hit = false; o.oob = false; o.point = s; o.ncode = 0; o.cell = 0;
    const uint32_t tiles = (uint32_t)f >> 3;
    const float hi = (float)f;
    const int ox = r.sx > 0, oy = r.sy > 0, oz = r.sz > 0;
    int steps = 0;
    for (int it = 0; it < kMaxSteps; ++it) {
        if (!dda_inside(q)) { o.oob = true; break; }
        const int qx = min(q.cx, f - 1), qy = min(q.cy, f - 1), qz = min(q.cz, f - 1);
        const uint32_t idx = FINE8 ? (uint32_t)(qx + (qy << 3) + (qz << 6)) : tiled_index(qx, qy, qz, tiles, tiles * tiles);
        if ((__ldg(mask + (idx >> 5)) >> (idx & 31u)) & 1u) {
            o.hit = true; o.cell = (uint32_t)qx | ((uint32_t)qy << 10) | ((uint32_t)qz << 20);
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
        o.ncode = mx ? (ox ? 1u : 2u) : (my ? (oy ? 3u : 4u) : (oz ? 5u : 6u));
    }
    o.steps = steps;
}

// The same march for the common case: 8^3 brick, no NaN among the tMax values.  Identical arithmetic; what changes is bookkeeping: one range test for the
// three cell coordinates, the bit index formed directly, the axis chosen by min (ties z over y over x, which is the
// reference's chain when nothing is NaN), crossing coordinates on the stepped axis taken from the cell faces.
__device__ __forceinline__ void fine_march8(const uint32_t* __restrict__ mask, V3 s, const Ray& r, FineOut& o) {
    int cx = (int)s.x, cy = (int)s.y, cz = (int)s.z;
    int ox = r.sx > 0, oy = r.sy > 0, oz = r.sz > 0;
    float tx = r.d.x != 0.0f ? fdiv(fsub((float)(cx + ox), s.x), r.d.x) : CUDART_INF_F;
    float ty = r.d.y != 0.0f ? fdiv(fsub((float)(cy + oy), s.y), r.d.y) : CUDART_INF_F;
    float tz = r.d.z != 0.0f ? fdiv(fsub((float)(cz + oz), s.z), r.d.z) : CUDART_INF_F;
    const bool on_edge = cx == 8 || cy == 8 || cz == 8;
    V3 point = s;
    uint32_t axis = 3u; // axis of the last counted step (3: none yet)
    int steps = 0;
    float dtx = r.dtx, dty = r.dty, dtz = r.dtz;
    int sx = r.sx, sy = r.sy, sz = r.sz;
#ifndef VX_FINE_NO_PIN
    // keep the loop invariants in registers: left alone, the compiler re-derives them from the direction in every round
    asm volatile("" : "+f"(dtx), "+f"(dty), "+f"(dtz), "+r"(sx), "+r"(sy), "+r"(sz), "+r"(ox), "+r"(oy), "+r"(oz), "+l"(mask));
#endif
    // The exits are labels so that the outcome flags are set once, after the loop, instead of being carried through it.
#pragma unroll 1
    for (;;) {
        uint32_t bit = (uint32_t)(cx + (cy << 3) + (cz << 6));
        if ((cx | cy | cz) & ~7) {
            // outside the brick, unless the march started on its far face (IsOnEdge, :216-232): then a coordinate equal to 8 is
            // inside on the axes the ray descends along, and the voxel looked up is the clamped one (:240-246)
            if (!on_edge || (unsigned)cx >= (unsigned)(ox ? 8 : 9) || (unsigned)cy >= (unsigned)(oy ? 8 : 9) || (unsigned)cz >= (unsigned)(oz ? 8 : 9))
                goto out_of_brick;
            bit = (uint32_t)(min(cx, 7) + (min(cy, 7) << 3) + (min(cz, 7) << 6));
        }
        if (cell_bit(mask, bit)) goto hit_voxel;
        float px, py, pz;
        uint32_t ax;
        // One step: axis of the minimum tMax (ties z over y over x), crossing point = start + t * direction except on the
        // stepped axis, where it is the face of the cell being left, (float)(cell + o); then tMax and the cell advance.
        asm("{\n\t.reg .pred qz, qy, qx;\n\t.reg .f32 m, t;\n\t.reg .s32 c;\n\t"
            "min.f32 m, %0, %1;\n\t"
            "setp.le.f32 qz, %2, m;\n\t"
            "setp.le.and.f32 qy|qx, %1, %0, !qz;\n\t"
            "selp.f32 t, %1, %0, qy;\n\t"
            "selp.f32 t, %2, t, qz;\n\t"
            "selp.u32 %9, 1, 0, qy;\n\t"
            "selp.u32 %9, 2, %9, qz;\n\t"
            "mul.rn.f32 %6, t, %13;\n\t"
            "mul.rn.f32 %7, t, %14;\n\t"
            "mul.rn.f32 %8, t, %15;\n\t"
            "add.rn.f32 %6, %10, %6;\n\t"
            "add.rn.f32 %7, %11, %7;\n\t"
            "add.rn.f32 %8, %12, %8;\n\t"
            "@qz add.s32 c, %5, %24;\n\t"
            "@qz cvt.rn.f32.s32 %8, c;\n\t"
            "@qz add.rn.f32 %2, %2, %18;\n\t"
            "@qz add.s32 %5, %5, %21;\n\t"
            "@qy add.s32 c, %4, %23;\n\t"
            "@qy cvt.rn.f32.s32 %7, c;\n\t"
            "@qy add.rn.f32 %1, %1, %17;\n\t"
            "@qy add.s32 %4, %4, %20;\n\t"
            "@qx add.s32 c, %3, %22;\n\t"
            "@qx cvt.rn.f32.s32 %6, c;\n\t"
            "@qx add.rn.f32 %0, %0, %16;\n\t"
            "@qx add.s32 %3, %3, %19;\n\t}"
            : "+f"(tx), "+f"(ty), "+f"(tz), "+r"(cx), "+r"(cy), "+r"(cz), "=f"(px), "=f"(py), "=f"(pz), "=r"(ax)
            : "f"(s.x), "f"(s.y), "f"(s.z), "f"(r.d.x), "f"(r.d.y), "f"(r.d.z), "f"(dtx), "f"(dty), "f"(dtz), "r"(sx), "r"(sy), "r"(sz),
              "r"(ox), "r"(oy), "r"(oz));
        if (fminf(fminf(px, py), pz) < 0.0f || fmaxf(fmaxf(px, py), pz) > 8.0f) goto out_of_brick;
        ++steps;
        point = mk(px, py, pz);
        axis = ax;
    }
hit_voxel:
    o.hit = true; o.oob = false; o.cell = (uint32_t)min(cx, 7) | ((uint32_t)min(cy, 7) << 10) | ((uint32_t)min(cz, 7) << 20);
    goto done;
out_of_brick:
    o.hit = false; o.oob = true; o.cell = 0;
done:
    o.point = point; o.steps = steps;
    o.ncode = axis == 3u ? 0u : (axis == 2u ? (oz ? 5u : 6u) : (axis == 1u ? (oy ? 3u : 4u) : (ox ? 1u : 2u)));
}

// ---------------------------------------------------------------------------------------------------------------------
// The brick march again, with everything that is not needed to FIND the end of the march taken out of the loop.
// What the reference's loop does per voxel (VolumeRaytracer.cu:236-349): range test, voxel lookup, axis choice, crossing
// point, test of the crossing point against [0,8]^3, step count, point and normal.  What decides where the march ends is
// only the lookup, the axis choice (the tMax values, accumulated with the same float additions) and the cell leaving the
// brick; the rest is recovered afterwards:
//   * the cell is a 9-bit voxel index plus one nibble per axis; leaving the brick sets bit 3 of a nibble (one AND tests all three);
//   * the step count is the sum of what the nibbles gained, the last axis is the nibble that moved last, and the crossing
//     point of the last counted step is rebuilt from its t (the smallest tMax of that round, which the loop keeps);
//   * the crossing-point test: on the stepped axis the coordinate is a voxel face (an integer in [0,8]: passes); on the
//     other axes it is fl(s + fl(t * d)), which is monotone in t, so over the whole march it stays between its values at
//     the first and the last crossing.  Those two are evaluated after the loop for all three axes (which is at least what
//     the reference tests); if any of the six lies outside [0,8] — a ray within rounding of a brick face, about one march
//     in a thousand — the function gives up and the caller runs the statement-by-statement march instead.
//   * a march that starts ON the far face of the brick (cell coordinate 8: IsOnEdge, :216-232 — every second entry through
//     a +x/+y/+z face, since the entry point is rounded) takes its first step out of the virtual cell before the loop when
//     that step is the first one the rule picks; every other combination (two such axes, another axis first) is left
//     to the caller's generic march.
// Returns false when the caller must run fine_march<true>() instead (nothing has been written to `o` then).
// 23 SASS per voxel step instead of 62.
// ---------------------------------------------------------------------------------------------------------------------
// The loop works in MIRRORED brick coordinates c' = (d > 0 ? c : 7 - c), in which every ray moves towards +x', +y', +z':
// the increments are the constants 1 / 8 / 64 (voxel index) and 1 / 16 / 256 (the per-axis cell nibbles, whose bit 3 comes
// on when a cell coordinate reaches 8, i.e. leaves the brick), whatever the direction, and the real voxel index is the
// mirrored one XOR a per-ray constant (7 - c == c ^ 7 on three bits).
constexpr uint32_t kFineOut = 0x888u;
#ifdef VX_FINE_STATS // diagnostics build (tools/build_variants.sh): how often the quick march hands over to the generic one, and why
__device__ unsigned long long g_fine_stats[8];
#define VX_FSTAT(i) atomicAdd(&g_fine_stats[i], 1ull)
#else
#define VX_FSTAT(i) ((void)0)
#endif

__device__ __forceinline__ void fine_step(float& tx, float& ty, float& tz, uint32_t& bit, uint32_t& u, float& t, float dtx, float dty, float dtz) {
    asm("{\n\t.reg .pred pz, py, px;\n\t"
        "min.f32 %5, %0, %1, %2;\n\t"                // t of this crossing: the smallest tMax
        "setp.le.f32 pz, %2, %5;\n\t"                // ties: z over y over x (:293-322 without NaNs)
        "setp.le.and.f32 py|px, %1, %5, !pz;\n\t"
        "@pz add.rn.f32 %2, %2, %8;\n\t"
        "@py add.rn.f32 %1, %1, %7;\n\t"
        "@px add.rn.f32 %0, %0, %6;\n\t"
        "@pz add.s32 %3, %3, 64;\n\t"
        "@py add.s32 %3, %3, 8;\n\t"
        "@px add.s32 %3, %3, 1;\n\t"
        "@pz add.s32 %4, %4, 256;\n\t"
        "@py add.s32 %4, %4, 16;\n\t"
        "@px add.s32 %4, %4, 1;\n\t}"
        : "+f"(tx), "+f"(ty), "+f"(tz), "+r"(bit), "+r"(u), "=f"(t)
        : "f"(dtx), "f"(dty), "f"(dtz));
}

template <bool SUN>
__device__ __forceinline__ bool fine_march8_quick(const uint32_t* __restrict__ mask, V3 s, const Ray& r, FineOut& o) {
    int cx = (int)s.x, cy = (int)s.y, cz = (int)s.z;
    if ((unsigned)cx > 8u || (unsigned)cy > 8u || (unsigned)cz > 8u) { VX_FSTAT(1); return false; }
    const int ox = r.sx > 0, oy = r.sy > 0, oz = r.sz > 0;
    float tx, ty, tz;
    if (VX_OPT_DPOS && SUN && r.dpos) {
        tx = fdiv(fsub((float)(cx + 1), s.x), r.d.x);
        ty = fdiv(fsub((float)(cy + 1), s.y), r.d.y);
        tz = fdiv(fsub((float)(cz + 1), s.z), r.d.z);
    } else if (r.dnz) {
        tx = tmax_div(fsub((float)(cx + ox), s.x), r.d.x);
        ty = tmax_div(fsub((float)(cy + oy), s.y), r.d.y);
        tz = tmax_div(fsub((float)(cz + oz), s.z), r.d.z);
    } else {
        tx = r.d.x != 0.0f ? fdiv(fsub((float)(cx + ox), s.x), r.d.x) : CUDART_INF_F;
        ty = r.d.y != 0.0f ? fdiv(fsub((float)(cy + oy), s.y), r.d.y) : CUDART_INF_F;
        tz = r.d.z != 0.0f ? fdiv(fsub((float)(cz + oz), s.z), r.d.z) : CUDART_INF_F;
    }
    VX_FSTAT(0);
    const float t_first = fminf(fminf(tx, ty), tz);
    int pre = 0; // steps taken before the loop (out of a virtual cell on the far face)
    const uint32_t edge = (cx == 8 ? 1u : 0u) | (cy == 8 ? 2u : 0u) | (cz == 8 ? 4u : 0u);
    if (edge) {
        // an axis that sits on the far face and does not descend: the cell is outside at once (:216-232, :283-287)
        if ((cx == 8 && !(r.d.x < 0.0f)) || (cy == 8 && !(r.d.y < 0.0f)) || (cz == 8 && !(r.d.z < 0.0f))) {
            o.hit = false; o.oob = true; o.cell = 0; o.point = s; o.steps = 0; o.ncode = 0;
            return true;
        }
        VX_FSTAT(5);
        if (edge & (edge - 1u)) { VX_FSTAT(2); return false; } // two axes on their far faces: generic march
        const bool pz = tz <= fminf(tx, ty), py = !pz && ty <= tx;
        const uint32_t first = pz ? 4u : (py ? 2u : 1u);
        if (first != edge) { VX_FSTAT(3); return false; }      // another axis moves first: generic march
        // the virtual cell looks up the clamped voxel (:240-246); then the first step leaves it through the face at 8
        cx = min(cx, 7); cy = min(cy, 7); cz = min(cz, 7);
        if (cell_bit(mask, (uint32_t)(cx + (cy << 3) + (cz << 6)))) {
            o.hit = true; o.oob = false; o.steps = 0; o.point = s; o.ncode = 0;
            o.cell = (uint32_t)cx | ((uint32_t)cy << 10) | ((uint32_t)cz << 20);
            return true;
        }
        if (edge == 1u) tx = fadd(tx, r.dtx); else if (edge == 2u) ty = fadd(ty, r.dty); else tz = fadd(tz, r.dtz);
        pre = 1;
    }
    const uint32_t mx = (uint32_t)(ox ? cx : 7 - cx), my = (uint32_t)(oy ? cy : 7 - cy), mz = (uint32_t)(oz ? cz : 7 - cz); // mirrored cell
    const uint32_t flip = (ox ? 0u : 7u) | (oy ? 0u : 56u) | (oz ? 0u : 448u);
    const uint32_t u0 = mx | (my << 4) | (mz << 8);
    uint32_t u = u0, up = u0;
    uint32_t bit = mx + (my << 3) + (mz << 6);
    float t_last = t_first, t_prev = t_first; // t of the last crossing and of the one before it
    float dtx = r.dtx, dty = r.dty, dtz = r.dtz;
    // keep the loop invariants in registers: left alone, ptxas re-derives the mask address and the deltas in every round
    asm volatile("" : "+f"(dtx), "+f"(dty), "+f"(dtz), "+l"(mask));
    bool hit;
#pragma unroll 1
    for (;;) {
        hit = cell_bit(mask, bit ^ flip);
        if (hit) break;
        up = u; t_prev = t_last;
        fine_step(tx, ty, tz, bit, u, t_last, dtx, dty, dtz);
        if (u & kFineOut) break;
    }
    const int steps = pre + (int)((u & 15u) + ((u >> 4) & 15u) + (u >> 8)) - (int)((u0 & 15u) + ((u0 >> 4) & 15u) + (u0 >> 8));
    if (steps == 0) { // the first voxel is solid
        o.hit = true; o.oob = false; o.steps = 0; o.point = s; o.ncode = 0;
        o.cell = (uint32_t)cx | ((uint32_t)cy << 10) | ((uint32_t)cz << 20);
        return true;
    }
    // the last counted step: the axis whose nibble moved last (the step out of the virtual cell when the loop took none)
    const uint32_t moved = u - up;
    const uint32_t axis = (u == u0) ? (edge == 1u ? 0u : (edge == 2u ? 1u : 2u)) : (moved == 1u ? 0u : (moved == 16u ? 1u : 2u));
    // crossing-point test (:325-341), see above.  Crossing k steps axis a_k at t_k and tests fl(s + fl(t_k * d)) on the other
    // two axes.  Checked here: the last crossing exactly as the reference tests it; the first crossing on all three axes
    // (when the march goes on after it, its stepped coordinate is an interior face, far from 0 and 8).  By monotonicity in t
    // that covers every crossing in between for the two axes that are not stepped last.  The axis stepped last is tested
    // by the reference only at earlier crossings, all at t <= t_prev: if its value at t_last is outside [0,8] (about every
    // second march that leaves the brick: the exit face, rounded) its value at t_prev bounds them instead.
    const float ax = fadd(s.x, fmul(t_first, r.d.x)), ay = fadd(s.y, fmul(t_first, r.d.y)), az = fadd(s.z, fmul(t_first, r.d.z));
    const float bx = fadd(s.x, fmul(t_last, r.d.x)), by = fadd(s.y, fmul(t_last, r.d.y)), bz = fadd(s.z, fmul(t_last, r.d.z));
    bool bad;
    {
        const float lo_b = fminf(axis == 0u ? 4.0f : bx, fminf(axis == 1u ? 4.0f : by, axis == 2u ? 4.0f : bz));
        const float hi_b = fmaxf(axis == 0u ? 4.0f : bx, fmaxf(axis == 1u ? 4.0f : by, axis == 2u ? 4.0f : bz));
        bad = lo_b < 0.0f || hi_b > 8.0f;
        if (steps > 1) {
            bad = bad || fminf(fminf(ax, ay), az) < 0.0f || fmaxf(fmaxf(ax, ay), az) > 8.0f;
            const float own = axis == 0u ? bx : (axis == 1u ? by : bz);
            if (own < 0.0f || own > 8.0f) {
                const float sa = axis == 0u ? s.x : (axis == 1u ? s.y : s.z), da = axis == 0u ? r.d.x : (axis == 1u ? r.d.y : r.d.z);
                const float v = fadd(sa, fmul(t_prev, da));
                bad = bad || v < 0.0f || v > 8.0f;
            }
        }
    }
    if (bad) { VX_FSTAT(4); return false; }
    // final cell: un-mirror the nibbles (8 = left the brick on that axis: cell 8 or -1)
    const int nx = (int)(u & 15u), ny = (int)((u >> 4) & 15u), nz = (int)(u >> 8);
    const int fx = ox ? nx : 7 - nx, fy = oy ? ny : 7 - ny, fz = oz ? nz : 7 - nz;
    // its crossing point: start + t * direction, except on the stepped axis, where it is the face of the cell that was left
    V3 p = mk(bx, by, bz);
    if (u == u0) { if (axis == 0u) p.x = 8.0f; else if (axis == 1u) p.y = 8.0f; else p.z = 8.0f; } // only the step out of the virtual cell
    else if (axis == 0u) p.x = (float)(ox ? fx : fx + 1);
    else if (axis == 1u) p.y = (float)(oy ? fy : fy + 1);
    else p.z = (float)(oz ? fz : fz + 1);
    o.hit = hit; o.oob = !hit; o.steps = steps; o.point = p;
    o.cell = hit ? ((uint32_t)fx | ((uint32_t)fy << 10) | ((uint32_t)fz << 20)) : 0u;
    o.ncode = axis == 2u ? (oz ? 5u : 6u) : (axis == 1u ? (oy ? 3u : 4u) : (ox ? 1u : 2u));
    return true;
}

// What Raytrace does with the outcome of one brick march (VolumeRaytracer.cu:409-523): a hit ends the ray (ST_DONE + F_HIT,
// result in pos/ncode); otherwise the start is re-seeded at the exit point with the reference's nudging rules (:438-489)
// and the lane begins a new coarse traversal.  (qx,qy,qz) = the brick, (cx,cy,cz) = the traversal cell it was found in.
template <bool BATCH>
__device__ __forceinline__ void brick_done(const WorldView& w, Lane& L, const FineOut& q, int qx, int qy, int qz, int cx, int cy, int cz,
                                           uint32_t next_axis, uint32_t hit_ncode, V3& pos, uint32_t& ncode) {
    const float ff = (float)w.factor;
    const float bx = fmul((float)qx, ff), by = fmul((float)qy, ff), bz = fmul((float)qz, ff);
    L.total += q.steps; L.fsteps += q.steps;
    const V3 hit_pos = mk(fadd(q.point.x, bx), fadd(q.point.y, by), fadd(q.point.z, bz));
    if (q.hit) {
        ncode = hit_ncode; // the slab test's normal if the brick march counted no step, else the last step's (:409-436)
        pos = hit_pos;
        if (L.total == 0) { // :515-523
            pos = mk(fmul(L.s.x, ff), fmul(L.s.y, ff), fmul(L.s.z, ff));
            ncode = (L.flags >> F_SN_SHIFT) & 7u;
        }
        if (BATCH) L.fcell = q.cell;
        L.flags |= F_HIT;
        lane_set_state(L, ST_DONE);
        return;
    }
    V3 start = to_brick_units(w, hit_pos);
    if (q.oob && qx == (int)start.x && qy == (int)start.y && qz == (int)start.z) { // :438-489
        start.x = next_toward(start.x, L.d.x < 0.0f);
        start.y = next_toward(start.y, L.d.y < 0.0f);
        start.z = next_toward(start.z, L.d.z < 0.0f);
        if (qx == (int)start.x && qy == (int)start.y && qz == (int)start.z) {
            // NextCell of the COARSE traversal = the cell after its final advance (:347, :473)
            const int sx = L.d.x > 0.0f ? 1 : -1, sy = L.d.y > 0.0f ? 1 : -1, sz = L.d.z > 0.0f ? 1 : -1;
            const int nx = cx + (next_axis == 0u ? sx : 0), ny = cy + (next_axis == 1u ? sy : 0), nz = cz + (next_axis == 2u ? sz : 0);
            const float gx = fsub((float)nx, start.x), gy = fsub((float)ny, start.y), gz = fsub((float)nz, start.z);
            const float ax = fabsf(gx), ay = fabsf(gy), az = fabsf(gz);
            if (ax < ay && ax < az) start.x = fadd(start.x, gx);
            else if (ay < ax && ay < az) start.y = fadd(start.y, gy);
            else start.z = fadd(start.z, gz);
        }
    }
    L.s = start;
    lane_set_state(L, ST_BEGIN);
}

// ST_CAND: the occupied cell L.bidx.  Tight-AABB slab test (VolumeRaytracer.cu:248-273); on a hit the brick is
// marched (:409-436) and the lane either finishes (ST_DONE + F_HIT, result in pos/ncode) or re-seeds its start
// (:438-489) and goes back to ST_BEGIN; on a slab miss it takes the step the reference takes and keeps going.
template <bool BATCH, bool COUNT, bool SUN>
__device__ __forceinline__ void lane_cand(const WorldView& w, Lane& L, V3& pos, uint32_t& ncode) {
    int qx, qy, qz;
    cell_of(w, L.bidx, qx, qy, qz);
    int cx = qx, cy = qy, cz = qz; // traversal cell (differs from the looked-up cell only on the padded far face)
    if (L.flags & F_SLOW) cell_of(w, L.idx, cx, cy, cz);
    else {
        if (!interior(w, qx, qy, qz)) { // border sentinel: the traversal left the grid (:283-287)
            const int it = steps_from_start(L, qx, qy, qz);
            L.total += it; L.csteps += it;
            lane_set_state(L, ST_DONE);
            return;
        }
        if (COUNT) { L.work.C++; L.work.K++; }
    }
    const uint2 rec = __ldg(w.precords + L.bidx);
    const Ray r = lane_ray(L);
    const uint32_t b = rec.x;
    const float l0 = (float)(b & 31u), l1 = (float)((b >> 5) & 31u), l2 = (float)((b >> 10) & 31u);
    const float h0 = (float)(((b >> 15) & 31u) + 1u), h1 = (float)(((b >> 20) & 31u) + 1u), h2 = (float)(((b >> 25) & 31u) + 1u);
    V3 lo, hi;
    if (VX_OPT_FMA && w.inv_factor >= 0.00390625f) {
        // factor = 2^k <= 256: l / factor + q = (l + q * factor) / factor with l <= 32 and q < 2^16 is a float, so neither the
        // product nor the sum rounds and one FMA gives the reference's two operations bit for bit
        const float inv_f = w.inv_factor;
        lo = mk(__fmaf_rn(l0, inv_f, (float)qx), __fmaf_rn(l1, inv_f, (float)qy), __fmaf_rn(l2, inv_f, (float)qz));
        hi = mk(__fmaf_rn(h0, inv_f, (float)qx), __fmaf_rn(h1, inv_f, (float)qy), __fmaf_rn(h2, inv_f, (float)qz));
    } else if (w.inv_factor != 0.0f) { // x / 2^k == x * 2^-k exactly
        const float inv_f = w.inv_factor;
        lo = mk(fadd(fmul(l0, inv_f), (float)qx), fadd(fmul(l1, inv_f), (float)qy), fadd(fmul(l2, inv_f), (float)qz));
        hi = mk(fadd(fmul(h0, inv_f), (float)qx), fadd(fmul(h1, inv_f), (float)qy), fadd(fmul(h2, inv_f), (float)qz));
    } else {
        const float fs = (float)w.factor;
        lo = mk(fadd(fdiv(l0, fs), (float)qx), fadd(fdiv(l1, fs), (float)qy), fadd(fdiv(l2, fs), (float)qz));
        hi = mk(fadd(fdiv(h0, fs), (float)qx), fadd(fdiv(h1, fs), (float)qy), fadd(fdiv(h2, fs), (float)qz));
    }
    const Slab sl = ray_aabb(L.s, r, lo, hi);
    if (!sl.hit) {
        if (VX_OPT_CSTEP && !(L.flags & F_SLOW)) { // fast mode: no NaN among the tMax values, so the fast loop's step is the reference's
            fast_step(L.tx, L.ty, L.tz, L.idx, r.dtx, r.dty, r.dtz, r.sx, L.d.y > 0.0f ? w.pnx : -w.pnx, L.d.z > 0.0f ? w.pnxy : -w.pnxy);
            lane_set_state(L, ST_FAST);
            return;
        }
        lane_step(w, L);
        lane_set_state(L, (L.flags & F_SLOW) ? ST_SLOW : ST_FAST);
        return;
    }
    const int it = steps_from_start(L, cx, cy, cz);
    const V3 point = it != 0 ? sl.point : L.s; // HitIntersectedPoint only moves `if (step != 0)` (:268-271)
    L.total += it; L.csteps += it;
    if (L.pc == L.bidx) { lane_set_state(L, ST_DONE); return; } // same brick as last time: give up (:402-407)
    L.pc = L.bidx;
    const float ff = (float)w.factor;
    const float bx = fmul((float)qx, ff), by = fmul((float)qy, ff), bz = fmul((float)qz, ff);
    const V3 local = mk(fsub(fmul(point.x, ff), bx), fsub(fmul(point.y, ff), by), fsub(fmul(point.z, ff), bz));
    if (COUNT) L.work.E++;
    // Axis of the coarse traversal's next advance (its NextCell, :347): the only thing the re-seeding below may still need
    // from the coarse state.  tMax and the cell index are dead from here on (every way out of this function that marches
    // again goes through lane_begin), and saying so keeps them out of registers during the brick march.
    const bool nmx = L.tx < L.ty && L.tx < L.tz;
    const bool nmy = !nmx && L.ty <= L.tx && L.ty < L.tz;
    const uint32_t next_axis = nmx ? 0u : (nmy ? 1u : 2u);
    L.tx = L.ty = L.tz = 0.0f; L.idx = 0;
    FineOut q;
    const uint32_t* mask = w.pool + (size_t)rec.y * w.words_per_brick;
    if (w.factor == 8) {
        // the tight variant whenever nothing can be NaN (the generic one keeps the reference's comparison order for NaNs)
#ifdef VX_FINE_OLD
        if (L.flags & F_CAPOK) fine_march8(mask, local, r, q);
        else fine_march<true>(mask, 8, local, r, q);
#else
        if (!((L.flags & F_CAPOK) && fine_march8_quick<SUN>(mask, local, r, q))) fine_march<true>(mask, 8, local, r, q);
#endif
    } else fine_march<false>(mask, w.factor, local, r, q);
    brick_done<BATCH>(w, L, q, qx, qy, qz, cx, cy, cz, next_axis, q.steps == 0 ? sl.ncode : q.ncode, pos, ncode);
}

// ---------------------------------------------------------------------------------------------
// Warp driver.  `Policy` supplies the rays and consumes the results:
//   uint64_t count()                      rays (ids) in this launch
//   int max_steps()
//   void prepare(L)                       once per thread before any ray (e.g. a direction common to all rays)
//   void init(w, id, L)                   set the lane up for ray `id` (lane_setup) or leave it ST_IDLE
//   void finish(w, L, pos, ncode)         consume a finished march and set the lane ST_IDLE (or re-arm it)
//   void flush()                          called by all 32 lanes whenever the whole warp is idle (before a refill / the end)
//   kRefill                               idle lanes that trigger a refill (32: whole tiles only)
//   kSun                                  all rays share one direction (see lane_begin)
//   kFastQuad, kLeaveNum, kLeaveDen       fast-loop scheduling, see lane_fast
//   kDrain, void drain(uint32_t done)     kDrain: called by all 32 lanes once per round, after finish(), with the ballot of the
//                                         lanes that finished this round (a policy batching its per-ray epilogue across the warp);
//                                         the same ballot is in pol.done_ballot while finish() runs
// `next` is a zero-initialised device counter shared by all warps of the launch.
// ---------------------------------------------------------------------------------------------
// ST_FAST: the hot loop over the bordered linear grid; leaving the grid shows up as an occupied (border) cell.
template <bool COUNT, int QUAD, int LEAVE_NUM, int LEAVE_DEN>
__device__ __forceinline__ void lane_fast(const WorldView& w, Lane& L) {
    static_assert(kFastChunk % QUAD == 0, "the step budget is counted in quads");
    if (L.flags & F_NEAR) { // rare (corner-to-corner views of the largest worlds): is a whole chunk still below the step cap?
        int x, y, z;
        cell_of(w, L.idx, x, y, z);
        if (steps_from_start(L, x, y, z) + kFastChunk > kMaxSteps) { L.flags |= F_SLOW; lane_set_state(L, ST_SLOW); return; }
    }
    const float dtx = lane_dt(L.d.x, L.ivx), dty = lane_dt(L.d.y, L.ivy), dtz = lane_dt(L.d.z, L.ivz);
    const int dix = L.d.x > 0.0f ? 1 : -1, diy = L.d.y > 0.0f ? w.pnx : -w.pnx, diz = L.d.z > 0.0f ? w.pnxy : -w.pnxy;
    float tx = L.tx, ty = L.ty, tz = L.tz;
    uint32_t idx = L.idx;
    int budget = kFastChunk;
    // Lanes leave this loop one by one as they meet occupied cells.  Once most of the lanes that entered it are
    // waiting, the few still stepping stop too (they stay ST_FAST and go on in the next round): a warp serves
    // its majority.  Checked every QUAD steps; a scheduling heuristic only, results do not depend on it.
    const int entered = __popc(__activemask());
    bool found = false;
#pragma unroll 1
    do {
#pragma unroll
        for (int k = 0; k < QUAD; ++k) {
            if (cell_bit(w.ptop, idx)) { found = true; break; }
            if (COUNT) L.work.C++;
            fast_step(tx, ty, tz, idx, dtx, dty, dtz, dix, diy, diz);
        }
        if (found) { lane_set_state(L, ST_CAND); break; }
        budget -= QUAD;
    } while (budget > 0 && __popc(__activemask()) * LEAVE_DEN >= entered * LEAVE_NUM);
    L.tx = tx; L.ty = ty; L.tz = tz; L.idx = idx; L.bidx = idx;
}

template <bool BATCH, bool COUNT, class Policy>
__device__ __forceinline__ void run_lanes(const WorldView& w, Policy& pol, unsigned long long* __restrict__ next) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t lt_mask = (1u << lane) - 1u;
    const unsigned long long count = pol.count();
    const int max_steps = pol.max_steps();
    Lane L;
    L.d = mk(0.0f, 0.0f, 0.0f); L.ivx = L.ivy = L.ivz = 0.0f; L.s = L.d; L.tx = L.ty = L.tz = 0.0f;
    L.idx = 0; L.bidx = 0; L.pc = kNoCell; L.flags = ST_IDLE; L.total = 0; L.csteps = L.fsteps = 0; L.fcell = 0;
    L.work.C = L.work.K = L.work.E = 0;
    pol.prepare(L); // policies whose rays share one direction set it here, once, so it never becomes per-lane state
    unsigned long long cur = 0, end = 0;
    bool exhausted = false;
    for (;;) {
        V3 pos = mk(0.0f, 0.0f, 0.0f);
        uint32_t ncode = 0;
        if (lane_state(L) == ST_BEGIN) lane_begin<Policy::kSun>(w, L, max_steps);
        if (lane_state(L) == ST_FAST) lane_fast<COUNT, Policy::kFastQuad, Policy::kLeaveNum, Policy::kLeaveDen>(w, L);
        if (lane_state(L) == ST_SLOW) lane_slow<COUNT>(w, L);
        if (lane_state(L) == ST_CAND) lane_cand<BATCH, COUNT, Policy::kSun>(w, L, pos, ncode);
        if constexpr (Policy::kDrain) {
            const uint32_t done = __ballot_sync(kFull, lane_state(L) == ST_DONE);
            pol.done_ballot = done; // finish() places its result by rank within this ballot, not by whoever happens to be converged
            if (lane_state(L) == ST_DONE) pol.finish(w, L, pos, ncode);
            if (done) pol.drain(done);
        } else {
            if (lane_state(L) == ST_DONE) pol.finish(w, L, pos, ncode);
        }
        const uint32_t idle = __ballot_sync(kFull, lane_state(L) == ST_IDLE);
        if (idle) {
            if (idle == kFull) pol.flush(); // the whole warp is between rays: results a policy holds back per tile go out now
            if (exhausted) { if (idle == kFull) break; continue; }
            if (__popc(idle) < Policy::kRefill) continue;
            if (cur == end) {
                unsigned long long base = 0;
                if (lane == 0) base = atomicAdd(next, (unsigned long long)kChunk);
                base = __shfl_sync(kFull, base, 0);
                if (base >= count) { exhausted = true; if (idle == kFull) break; continue; }
                cur = base; end = base + kChunk < count ? base + kChunk : count;
            }
            const uint32_t avail = (uint32_t)(end - cur), want = (uint32_t)__popc(idle);
            const uint32_t take = want < avail ? want : avail;
            const uint32_t rank = (uint32_t)__popc(idle & lt_mask);
            if (lane_state(L) == ST_IDLE && rank < take) pol.init(w, cur + rank, L);
            cur += take;
        }
    }
}

} // namespace vx
