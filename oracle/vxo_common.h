/*
 * vxo_common.h — shared helpers of the CPU ORACLE (test infrastructure, not product code).
 * Scalar helpers that pin down the arithmetic the reference's device code relies on.
 */
#ifndef VXO_COMMON_H
#define VXO_COMMON_H
#include <cmath>
#include <cstdint>
#include <cstring>
#include <functional>
#include <limits>
#include <thread>
#include <vector>

namespace vxo {

constexpr float kInf = std::numeric_limits<float>::infinity();
constexpr float kFltEps = std::numeric_limits<float>::epsilon(); /* VolumeRaytracer.cuh:22 */
constexpr double kEpsDDA = 1e-6;                                 /* VolumeRaytracer.cuh:20 (a double) */
constexpr int kMaxSteps = 2048;                                  /* VolumeRaytracer.cuh:235 */

struct V3 { float x, y, z; };

/* CUDA cvt.rzi.s32.f32: truncation, saturating, NaN -> 0 */
static inline int f2i(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return 2147483647;
    if (v <= -2147483648.0f) return -2147483647 - 1;
    return (int)v;
}
/* CUDA cvt.rzi.u32.f32: truncation, saturating, NaN -> 0, negative -> 0 */
static inline uint32_t f2u_cuda(float v) {
    if (v != v) return 0u;
    if (v <= 0.0f) return 0u;
    if (v >= 4294967296.0f) return 0xFFFFFFFFu;
    return (uint32_t)v;
}
/* what x86-64 g++ emits for (unsigned)(float): cvttss2si r64, low 32 bits */
static inline uint32_t f2u_x86(float v) {
    if (v != v || v >= 9223372036854775808.0f || v < -9223372036854775808.0f) return 0u;
    return (uint32_t)(uint64_t)(int64_t)v;
}
/* CUDA fminf/fmaxf: NaN-suppressing, -0 < +0 */
static inline float fmin_cu(float a, float b) {
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) return std::signbit(a) ? a : b;
    return a < b ? a : b;
}
static inline float fmax_cu(float a, float b) {
    if (a != a) return b;
    if (b != b) return a;
    if (a == b) return std::signbit(a) ? b : a;
    return a > b ? a : b;
}

/* helper_math.h:1325-1329 with the HOST rsqrtf of CUDA's crt/math_functions.hpp:3333-3336 */
static inline V3 normalize_host(V3 v) {
    float d = v.x * v.x + v.y * v.y + v.z * v.z;
    float inv = (float)(1.0 / std::sqrt((double)d));
    return V3{v.x * inv, v.y * inv, v.z * inv};
}

/* tiled-linear index, VolumeRaytracer.cuh:111-131, widened to 64 bit */
static inline uint64_t sample_index(uint32_t x, uint32_t y, uint32_t z, uint32_t width, uint32_t height) {
    const uint64_t tW = width / 8u, tH = height / 8u;
    const uint64_t tile = (x / 8u) + (y / 8u) * tW + (uint64_t)(z / 8u) * tW * tH;
    const uint32_t fine = (x % 8u) + (y % 8u) * 8u + (z % 8u) * 64u;
    return tile * 512u + fine;
}
static inline bool bit_get(const uint32_t* words, uint64_t i) { return (words[i >> 5] >> (i & 31u)) & 1u; }
static inline void bit_set(uint32_t* words, uint64_t i) { words[i >> 5] |= 1u << (i & 31u); }

static inline void parallel_for(size_t n, int nthreads, const std::function<void(size_t, size_t)>& body) {
    if (nthreads <= 0) nthreads = (int)std::thread::hardware_concurrency();
    if (nthreads < 1) nthreads = 1;
    if ((size_t)nthreads > n) nthreads = n ? (int)n : 1;
    if (nthreads == 1) { body(0, n); return; }
    std::vector<std::thread> pool;
    /* many small chunks dealt round-robin: terrain cost varies strongly across rows */
    const size_t chunks = (size_t)nthreads * 16;
    const size_t per = (n + chunks - 1) / chunks;
    for (int t = 0; t < nthreads; ++t)
        pool.emplace_back([=, &body]() {
            for (size_t c = (size_t)t; c * per < n; c += (size_t)nthreads) {
                size_t lo = c * per, hi = lo + per < n ? lo + per : n;
                body(lo, hi);
            }
        });
    for (auto& th : pool) th.join();
}

} // namespace vxo
#endif
