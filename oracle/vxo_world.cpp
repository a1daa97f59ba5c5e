/*
 * vxo_world.cpp — CPU ORACLE (test infrastructure, not product code): noise terrain + brickmap build.
 *
 * Restates, for the CPU:
 *   cuda_noise.cuh:44-54 (hash), :118-121 (randomIntGrid), :162-165 (lerp), :174-196 (grad),
 *   :199-202 (fade), :565-610 (perlinNoise), :615-628 (repeaterPerlin);
 *   VoxelWorldBuilder.cu:4-35 (PerlinNoise, PopulateVoxels);
 *   VolumeRaytracer.cuh:379-516 (GenerateLowresVoxelBuffer).
 * Parity status: pinned for the brickmap build against the reference's own
 * GenerateLowresVoxelBuffer compiled for the host (oracle/_ref, tests/test_oracle.py, tests/test_reference_pins.py);
 * the noise chain is pinned against the reference's cuda_noise.cuh compiled as plain C++
 * (x86 float->uint mode) and, on the GPU box, against the reference's PopulateVoxels kernel.
 */
#include "vxo.h"
#include "vxo_common.h"
#include "vxo_world.h"

using namespace vxo;

/* ---- noise ---- */
extern "C" uint32_t vxo_hash(uint32_t s) {
    s = (s + 0x7ed55d16u) + (s << 12);
    s = (s ^ 0xc761c23cu) ^ (s >> 19);
    s = (s + 0x165667b1u) + (s << 5);
    s = (s + 0xd3a2646cu) ^ (s << 9);
    s = (s + 0xfd7046c5u) + (s << 3);
    s = (s ^ 0xb55a4f09u) ^ (s >> 16);
    return s;
}

static inline uint32_t lattice_hash(float x, float y, float z, float seed, int mode) {
    /* ((x*1723 + y*93241) + z*149812) + 3824 + seed, all binary32, then float->uint */
    float a = x * 1723.0f;
    float b = y * 93241.0f;
    float c = z * 149812.0f;
    float s = a + b;
    s = s + c;
    s = s + 3824.0f;
    s = s + seed;
    return vxo_hash(mode == VXO_F2U_X86 ? f2u_x86(s) : f2u_cuda(s));
}

static inline float grad16(uint32_t h, float x, float y, float z) {
    switch (h & 0xFu) {
    case 0x0: return x + y;
    case 0x1: return -x + y;
    case 0x2: return x - y;
    case 0x3: return -x - y;
    case 0x4: return x + z;
    case 0x5: return -x + z;
    case 0x6: return x - z;
    case 0x7: return -x - z;
    case 0x8: return y + z;
    case 0x9: return -y + z;
    case 0xA: return y - z;
    case 0xB: return -y - z;
    case 0xC: return y + x;
    case 0xD: return -y + z;
    case 0xE: return y - x;
    default:  return -y - z;
    }
}
static inline float fade5(float t) {
    float inner = t * 6.0f - 15.0f;
    inner = t * inner + 10.0f;
    return t * t * t * inner;
}
static inline float mix(float a, float b, float r) {
    float wa = 1.0f - r;
    return a * wa + b * r;
}

extern "C" float vxo_perlin(float px, float py, float pz, float scale, int seed, int mode) {
    const float fseed = (float)seed;
    px = px * scale; py = py * scale; pz = pz * scale;
    const float ix = std::floor(px), iy = std::floor(py), iz = std::floor(pz);
    px -= ix; py -= iy; pz -= iz;
    const float u = fade5(px), v = fade5(py), w = fade5(pz);
    const float ix1 = ix + 1.0f, iy1 = iy + 1.0f, iz1 = iz + 1.0f;
    const float qx = px - 1.0f, qy = py - 1.0f, qz = pz - 1.0f;
    const float i000 = grad16(lattice_hash(ix,  iy,  iz,  fseed, mode), px, py, pz);
    const float i100 = grad16(lattice_hash(ix1, iy,  iz,  fseed, mode), qx, py, pz);
    const float i010 = grad16(lattice_hash(ix,  iy1, iz,  fseed, mode), px, qy, pz);
    const float i110 = grad16(lattice_hash(ix1, iy1, iz,  fseed, mode), qx, qy, pz);
    const float i001 = grad16(lattice_hash(ix,  iy,  iz1, fseed, mode), px, py, qz);
    const float i101 = grad16(lattice_hash(ix1, iy,  iz1, fseed, mode), qx, py, qz);
    const float i011 = grad16(lattice_hash(ix,  iy1, iz1, fseed, mode), px, qy, qz);
    const float i111 = grad16(lattice_hash(ix1, iy1, iz1, fseed, mode), qx, qy, qz);
    const float x00 = mix(i000, i100, u), x10 = mix(i010, i110, u);
    const float x01 = mix(i001, i101, u), x11 = mix(i011, i111, u);
    const float y0 = mix(x00, x10, v), y1 = mix(x01, x11, v);
    return mix(y0, y1, w);
}

extern "C" float vxo_terrain_noise(float fx, float fy, float fz, int mode) {
    /* repeaterPerlin(pos, 1.0f, <ignored seed>, 32, 2.0f, 0.5f) — VoxelWorldBuilder.cu:6 */
    float acc = 0.0f, amp = 1.0f, scale = 1.0f;
    for (int i = 0; i < 32; ++i) {
        float n = vxo_perlin(fx * scale, fy * scale, fz * scale, 1.0f, (i + 38) * 27389482, mode);
        acc += n * amp;
        scale *= 2.0f;
        amp *= 0.5f;
    }
    return acc;
}

extern "C" int vxo_voxel_solid(uint32_t x, uint32_t y, uint32_t z, int mode) {
    const float scale = 0.005f; /* float scale = 0.005; VoxelWorldBuilder.cu:17 */
    float t = vxo_terrain_noise((float)(int)x * scale, (float)(int)y * scale, (float)(int)z * scale, mode) * 1000.0f;
    t = fmax_cu(t, 0.0f);
    return !((float)(int)y > t);
}

extern "C" void vxo_populate_dense(uint32_t* bits, uint32_t X, uint32_t Y, uint32_t Z,
                                   uint32_t x0, uint32_t y0, uint32_t z0, int mode, int nthreads) {
    const uint64_t nvox = (uint64_t)X * Y * Z;
    std::memset(bits, 0, (size_t)((nvox + 31) / 32) * 4);
    const uint32_t tX = X / 8, tY = Y / 8, tZ = Z / 8;
    parallel_for((size_t)tX * tY * tZ, nthreads, [&](size_t lo, size_t hi) {
        for (size_t t = lo; t < hi; ++t) {
            uint32_t bx = (uint32_t)(t % tX), by = (uint32_t)((t / tX) % tY), bz = (uint32_t)(t / ((size_t)tX * tY));
            for (uint32_t f = 0; f < 512; ++f) {
                uint32_t x = bx * 8 + (f & 7), y = by * 8 + ((f >> 3) & 7), z = bz * 8 + (f >> 6);
                if (vxo_voxel_solid(x + x0, y + y0, z + z0, mode)) bit_set(bits, (uint64_t)t * 512 + f);
            }
        }
    });
}

/* ---- brickmap ---- */
extern "C" uint64_t vxo_sample_index(uint32_t x, uint32_t y, uint32_t z, uint32_t w, uint32_t h) {
    return sample_index(x, y, z, w, h);
}
extern "C" void vxo_position_from_index(uint64_t index, uint32_t width, uint32_t height,
                                        uint32_t* x, uint32_t* y, uint32_t* z) {
    /* VolumeRaytracer.cuh:143-163 */
    const uint64_t tW = width / 8u, tH = height / 8u;
    const uint64_t tile = index / 512u;
    const uint32_t fine = (uint32_t)(index % 512u);
    *x = (uint32_t)(tile % tW) * 8u + fine % 8u;
    *y = (uint32_t)((tile / tW) % tH) * 8u + (fine / 8u) % 8u;
    *z = (uint32_t)(tile / (tW * tH)) * 8u + fine / 64u;
}

static void build_bricks(vxo_world* w, int nthreads) {
    const uint32_t f = (uint32_t)w->factor;
    const uint32_t cx = w->X / f, cy = w->Y / f, cz = w->Z / f;
    w->cx = cx; w->cy = cy; w->cz = cz;
    const size_t nb = (size_t)cx * cy * cz;
    const size_t wpb = (size_t)f * f * f / 32;
    w->words_per_brick = wpb;
    w->brick_bits.assign(nb * wpb, 0u);
    w->bounds.assign(nb * 6, 0.0f);
    std::vector<uint8_t> any(nb, 0);
    parallel_for(nb, nthreads, [&](size_t lo, size_t hi) {
        for (size_t t = lo; t < hi; ++t) {
            uint32_t bx, by, bz;
            vxo_position_from_index(t, cx, cy, &bx, &by, &bz);
            const size_t idx = (size_t)sample_index(bx, by, bz, cx, cy); /* == t when cx,cy are multiples of 8 */
            uint32_t* mask = &w->brick_bits[idx * wpb];
            int mn[3] = {INT32_MAX, INT32_MAX, INT32_MAX}, mx[3] = {INT32_MIN, INT32_MIN, INT32_MIN};
            bool occ = false;
            for (uint32_t dz = 0; dz < f; ++dz)
                for (uint32_t dy = 0; dy < f; ++dy)
                    for (uint32_t dx = 0; dx < f; ++dx) {
                        const uint64_t hi_idx = sample_index(dx + f * bx, dy + f * by, dz + f * bz, w->X, w->Y);
                        if (!bit_get(w->dense.data(), hi_idx)) continue;
                        bit_set(mask, sample_index(dx, dy, dz, f, f));
                        occ = true;
                        mn[0] = std::min(mn[0], (int)dx); mx[0] = std::max(mx[0], (int)dx);
                        mn[1] = std::min(mn[1], (int)dy); mx[1] = std::max(mx[1], (int)dy);
                        mn[2] = std::min(mn[2], (int)dz); mx[2] = std::max(mx[2], (int)dz);
                    }
            if (!occ) { mn[0] = mn[1] = mn[2] = 0; mx[0] = mx[1] = mx[2] = -1; }
            float* b = &w->bounds[idx * 6];
            for (int k = 0; k < 3; ++k) { b[k] = (float)mn[k]; b[3 + k] = (float)mx[k]; }
            any[idx] = occ;
        }
    });
    w->top.assign((nb + 31) / 32, 0u);
    for (size_t i = 0; i < nb; ++i) if (any[i]) bit_set(w->top.data(), i);
    w->bind();
}

/* Adopt a brickmap that already exists in the reference's form (e.g. downloaded from the GPU build
 * for the timed CPU baseline at sizes the CPU cannot generate): no copy, the caller keeps the
 * arrays alive.  There is no dense grid in such a world. */
extern "C" vxo_world* vxo_world_from_brickmap_view(const uint32_t* top, const uint32_t* brick_bits, const float* bounds6,
                                                   uint32_t X, uint32_t Y, uint32_t Z, int factor) {
    if (factor < 8 || factor % 8 || X % factor || Y % factor || Z % factor) return nullptr;
    vxo_world* w = new vxo_world();
    w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    w->cx = X / factor; w->cy = Y / factor; w->cz = Z / factor;
    w->words_per_brick = (size_t)factor * factor * factor / 32;
    w->top_p = top; w->brick_p = brick_bits; w->bounds_p = bounds6;
    return w;
}

extern "C" vxo_world* vxo_world_from_dense(const uint32_t* bits, uint32_t X, uint32_t Y, uint32_t Z, int factor) {
    if (factor < 8 || factor % 8 || X % factor || Y % factor || Z % factor) return nullptr;
    /* the coarse grid is addressed with the same 8^3-tiled index */
    if ((X / factor) % 8 || (Y / factor) % 8 || (Z / factor) % 8) return nullptr;
    vxo_world* w = new vxo_world();
    w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    const uint64_t nvox = (uint64_t)X * Y * Z;
    w->dense.assign(bits, bits + (size_t)((nvox + 31) / 32));
    build_bricks(w, 0);
    return w;
}

extern "C" vxo_world* vxo_world_from_noise(uint32_t X, uint32_t Y, uint32_t Z, int factor, int mode, int nthreads) {
    if (factor < 8 || factor % 8 || X % factor || Y % factor || Z % factor) return nullptr;
    if ((X / factor) % 8 || (Y / factor) % 8 || (Z / factor) % 8) return nullptr;
    vxo_world* w = new vxo_world();
    w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    const uint64_t nvox = (uint64_t)X * Y * Z;
    w->dense.resize((size_t)((nvox + 31) / 32));
    vxo_populate_dense(w->dense.data(), X, Y, Z, 0, 0, 0, mode, nthreads);
    build_bricks(w, nthreads);
    return w;
}

extern "C" void vxo_world_free(vxo_world* w) { delete w; }
extern "C" void vxo_world_dims(const vxo_world* w, uint32_t d[3], uint32_t c[3], int* factor) {
    d[0] = w->X; d[1] = w->Y; d[2] = w->Z; c[0] = w->cx; c[1] = w->cy; c[2] = w->cz; *factor = w->factor;
}
extern "C" const uint32_t* vxo_world_dense_bits(const vxo_world* w) { return w->dense.data(); }
extern "C" const uint32_t* vxo_world_top_bits(const vxo_world* w) { return w->top_p; }
extern "C" const uint32_t* vxo_world_brick_bits(const vxo_world* w) { return w->brick_p; }
extern "C" const float* vxo_world_bounds(const vxo_world* w) { return w->bounds_p; }
