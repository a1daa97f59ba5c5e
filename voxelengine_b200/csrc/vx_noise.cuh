// vx_noise.cuh — the terrain density function of the reference's world builder, bit-exact.
//
// solid(x,y,z) = !(y > max(1000 * fBm(0.005*(x,y,z)), 0))      VoxelWorldBuilder.cu:4-35
// fBm = 32 octaves of Perlin gradient noise, lacunarity 2, decay 0.5, seed_i = (i+38)*27389482
//                                                              cuda_noise.cuh:565-628
// The float->unsigned conversion inside the lattice hash is CUDA's saturating cvt.rzi.u32.f32
// (cuda_noise.cuh:118-121).  Arithmetic is IEEE binary32 without contraction.
//
// New here: an exact early-out.  After k octaves the remaining octaves can move the sum by at
// most R_k = B*2^(1-k) + E (B bounds |perlin|, E the float accumulation error), so the comparison
// y > t is already decided for every voxel farther than 1000*R_k from the iso-surface.  Voxels
// inside that band run all 32 octaves, which keeps the result identical bit for bit.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vx {

__device__ __forceinline__ uint32_t hash32(uint32_t s) { // cuda_noise.cuh:44-54
    s = (s + 0x7ed55d16u) + (s << 12);
    s = (s ^ 0xc761c23cu) ^ (s >> 19);
    s = (s + 0x165667b1u) + (s << 5);
    s = (s + 0xd3a2646cu) ^ (s << 9);
    s = (s + 0xfd7046c5u) + (s << 3);
    s = (s ^ 0xb55a4f09u) ^ (s >> 16);
    return s;
}

__device__ __forceinline__ float grad16(uint32_t h, float x, float y, float z) { // cuda_noise.cuh:174-196
    // 16 two-term sums +-a +-b: rows 0-7 lead with x, rows 8-15 with y; second term y (0-3),
    // z (4-11), then x,z,x,z (12-15); bit 0 negates the first term, bit 1 the second.
    h &= 15u;
    float a = h < 8u ? x : y;
    float b = h < 4u ? y : (h < 12u ? z : ((h & 1u) ? z : x));
    if (h & 1u) a = -a;
    if (h & 2u) b = -b;
    return __fadd_rn(a, b);
}

__device__ __forceinline__ float fade5(float t) { // cuda_noise.cuh:199-202
    float inner = __fsub_rn(__fmul_rn(t, 6.0f), 15.0f);
    inner = __fadd_rn(__fmul_rn(t, inner), 10.0f);
    return __fmul_rn(__fmul_rn(__fmul_rn(t, t), t), inner);
}
__device__ __forceinline__ float mixf(float a, float b, float r) { // cuda_noise.cuh:162-165
    return __fadd_rn(__fmul_rn(a, __fsub_rn(1.0f, r)), __fmul_rn(b, r));
}

// perlinNoise(pos, 1.0f, seed) — cuda_noise.cuh:565-610
__device__ __forceinline__ float perlin3(float px, float py, float pz, float fseed) {
    const float ix = floorf(px), iy = floorf(py), iz = floorf(pz);
    px = __fsub_rn(px, ix); py = __fsub_rn(py, iy); pz = __fsub_rn(pz, iz);
    const float u = fade5(px), v = fade5(py), w = fade5(pz);
    const float qx = __fsub_rn(px, 1.0f), qy = __fsub_rn(py, 1.0f), qz = __fsub_rn(pz, 1.0f);
    const float a0 = __fmul_rn(ix, 1723.0f), a1 = __fmul_rn(__fadd_rn(ix, 1.0f), 1723.0f);
    const float b0 = __fmul_rn(iy, 93241.0f), b1 = __fmul_rn(__fadd_rn(iy, 1.0f), 93241.0f);
    const float c0 = __fmul_rn(iz, 149812.0f), c1 = __fmul_rn(__fadd_rn(iz, 1.0f), 149812.0f);
    const float ab00 = __fadd_rn(a0, b0), ab10 = __fadd_rn(a1, b0), ab01 = __fadd_rn(a0, b1), ab11 = __fadd_rn(a1, b1);
#define VX_LATTICE(ab, c) hash32((unsigned int)__fadd_rn(__fadd_rn(__fadd_rn(ab, c), 3824.0f), fseed))
    const float i000 = grad16(VX_LATTICE(ab00, c0), px, py, pz);
    const float i100 = grad16(VX_LATTICE(ab10, c0), qx, py, pz);
    const float i010 = grad16(VX_LATTICE(ab01, c0), px, qy, pz);
    const float i110 = grad16(VX_LATTICE(ab11, c0), qx, qy, pz);
    const float i001 = grad16(VX_LATTICE(ab00, c1), px, py, qz);
    const float i101 = grad16(VX_LATTICE(ab10, c1), qx, py, qz);
    const float i011 = grad16(VX_LATTICE(ab01, c1), px, qy, qz);
    const float i111 = grad16(VX_LATTICE(ab11, c1), qx, qy, qz);
#undef VX_LATTICE
    const float x00 = mixf(i000, i100, u), x10 = mixf(i010, i110, u);
    const float x01 = mixf(i001, i101, u), x11 = mixf(i011, i111, u);
    return mixf(mixf(x00, x10, v), mixf(x01, x11, v), w);
}

constexpr int kOctaves = 32;
// |perlin| <= 2 (gradients are sums of two terms in [-1,1]; the three lerps are convex up to
// rounding); 2.001 leaves 5e-4 relative slack.  Accumulation error of <= 32 float adds of partial
// sums below 4.01: 32 * 2^-24 * 4.01 < 7.7e-6.
constexpr double kPerlinBound = 2.001;
constexpr double kAccumSlack = 7.7e-6;

// State of one voxel's evaluation, advanced one octave at a time so that a warp can keep all
// lanes on the same instruction while they work on different voxels/octaves.
struct Terrain {
    float fx, fy, fz, yv;
    float acc, amp, scale;
    int k;
    __device__ __forceinline__ void begin(int x, int y, int z) {
        const float s = 0.005f; // float scale = 0.005; VoxelWorldBuilder.cu:17
        fx = __fmul_rn((float)x, s); fy = __fmul_rn((float)y, s); fz = __fmul_rn((float)z, s);
        yv = (float)y;
        acc = 0.0f; amp = 1.0f; scale = 1.0f; k = 0;
    }
    __device__ __forceinline__ void octave() {
        const float n = perlin3(__fmul_rn(fx, scale), __fmul_rn(fy, scale), __fmul_rn(fz, scale),
                                (float)((k + 38) * 27389482));
        acc = __fadd_rn(acc, __fmul_rn(n, amp));
        scale = __fmul_rn(scale, 2.0f);
        amp = __fmul_rn(amp, 0.5f);
        ++k;
    }
    // -1: undecided, 0: empty, 1: solid.  Exact (see header comment).
    __device__ __forceinline__ int decided() const {
        if (k >= kOctaves) {
            const float t = fmaxf(__fmul_rn(acc, 1000.0f), 0.0f);
            return !(yv > t);
        }
        if (yv == 0.0f) return 1; // t >= 0 always, so !(0 > t)
        const double r = kPerlinBound * 2.0 * (double)amp + kAccumSlack; // amp == 2^-k exactly
        const double lo = ((double)acc - r) * 1000.0, hi = ((double)acc + r) * 1000.0;
        const double y = (double)yv;
        if (lo > 0.0 && lo * (1.0 - 1.2e-7) >= y) return 1;
        const double top = hi > 0.0 ? hi * (1.0 + 1.2e-7) : 0.0;
        if (top < y) return 0;
        return -1;
    }
};

__device__ __forceinline__ bool voxel_solid_full(int x, int y, int z) {
    Terrain t; t.begin(x, y, z);
    for (int i = 0; i < kOctaves; ++i) t.octave();
    return t.decided() == 1;
}

} // namespace vx
