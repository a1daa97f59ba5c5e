// vx_world.cu — brickmap construction on the device.
//
// Replaces the reference's start-up chain: PopulateVoxels kernel + dense D2H copy
// (VoxelWorldBuilder.cuh:12-32), the host-thread GenerateLowresVoxelBuffer (VolumeRaytracer.cuh:379-516)
// and the per-brick cudaMalloc uploads (VolumeRaytracer.cu:527-572) by four kernels:
//   A  noise_tiles / dense_tiles : one warp per 8^3 tile, ballot-assembled 512-bit masks
//   B  analyse_bricks            : occupancy, packed tight bounds, voxel count (one warp per brick)
//   C  scan (3 small kernels)    : exclusive prefix over occupancy -> pool slot
//   D  compact_bricks            : copy occupied bricks into the compact pool
// Layout: DESIGN.md §2.
#include "vx_host.h"
#include "vx_noise.cuh"

namespace vx {

// ---- brick/tile enumeration ---------------------------------------------------------------
// Global tile id t = brick * tiles_per_brick + sub, brick = reference tiled coarse index.
struct BuildGeom {
    uint32_t cx, cy, cz;     // coarse dims
    uint32_t f8;             // factor / 8
    uint32_t tpb;            // tiles per brick = f8^3
    uint32_t x0, y0, z0;     // noise window origin
    uint32_t X, Y;           // voxel dims (dense source)
};
__device__ __forceinline__ void brick_coords(uint32_t brick, const BuildGeom& g, uint32_t& bx, uint32_t& by, uint32_t& bz) {
    // inverse of the tiled index, VolumeRaytracer.cuh:143-163
    const uint32_t tW = g.cx >> 3, tH = g.cy >> 3;
    const uint32_t tile = brick >> 9, fine = brick & 511u;
    bx = (tile % tW) * 8u + (fine & 7u);
    by = ((tile / tW) % tH) * 8u + ((fine >> 3) & 7u);
    bz = (tile / (tW * tH)) * 8u + (fine >> 6);
}

// ---- A1: noise -> tile masks ----------------------------------------------------------------
// One warp per 8^3 tile.  Lane l owns the 16 voxels {i*32 + l}; a flattened (voxel, octave) loop
// keeps the warp converged while lanes sit in different octaves (early-out, vx_noise.cuh).
template <bool EARLY_OUT>
__global__ void __launch_bounds__(256) noise_tiles(uint32_t* __restrict__ tmp, BuildGeom g, uint64_t ntiles, uint64_t tile0) {
    // tiles tile0 .. tile0 + ntiles - 1 of the world (a z-slab of a partitioned build; the whole world with tile0 = 0) into tmp[0..]
    const uint32_t lane = threadIdx.x & 31u;
    const uint64_t warps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    for (uint64_t t = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5; t < ntiles; t += warps) {
        const uint32_t brick = (uint32_t)((t + tile0) / g.tpb), sub = (uint32_t)((t + tile0) % g.tpb);
        uint32_t bx, by, bz;
        brick_coords(brick, g, bx, by, bz);
        const uint32_t f = g.f8 * 8u;
        const int vx0 = (int)(g.x0 + bx * f + (sub % g.f8) * 8u + (lane & 7u));
        const int vy0 = (int)(g.y0 + by * f + ((sub / g.f8) % g.f8) * 8u + (lane >> 3));
        const int vz0 = (int)(g.z0 + bz * f + (sub / (g.f8 * g.f8)) * 8u);
        uint32_t bits = 0; // bit i = voxel i*32+lane: y = vy0 + 4*(i&1), z = vz0 + (i>>1)
        int i = 0;
        Terrain tr;
        tr.begin(vx0, vy0, vz0);
        while (i < 16) {
            tr.octave();
            int dec;
            if (EARLY_OUT) dec = tr.decided();
            else dec = tr.k >= kOctaves ? tr.decided() : -1;
            if (dec >= 0) {
                bits |= (uint32_t)dec << i;
                ++i;
                tr.begin(vx0, vy0 + ((i & 1) << 2), vz0 + (i >> 1));
            }
        }
        uint32_t mine = 0;
#pragma unroll
        for (int w = 0; w < 16; ++w) {
            const uint32_t word = __ballot_sync(0xFFFFFFFFu, (bits >> w) & 1u);
            if (lane == (uint32_t)w) mine = word;
        }
        if (lane < 16u) tmp[t * 16u + lane] = mine;
    }
}

// ---- A2: external dense grid (reference tiled layout) -> tile masks --------------------------
__global__ void __launch_bounds__(256) dense_tiles(uint32_t* __restrict__ tmp, const uint32_t* __restrict__ dense, BuildGeom g,
                                                   uint64_t ntiles) {
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; // one thread per 32-bit word
    if (gid >= ntiles * 16u) return;
    const uint64_t t = gid >> 4;
    const uint32_t wd = (uint32_t)gid & 15u;
    const uint32_t brick = (uint32_t)(t / g.tpb), sub = (uint32_t)(t % g.tpb);
    uint32_t bx, by, bz;
    brick_coords(brick, g, bx, by, bz);
    const uint64_t tx = (uint64_t)bx * g.f8 + sub % g.f8, ty = (uint64_t)by * g.f8 + (sub / g.f8) % g.f8,
                   tz = (uint64_t)bz * g.f8 + sub / (g.f8 * g.f8);
    const uint64_t src_tile = tx + ty * (g.X >> 3) + tz * (uint64_t)(g.X >> 3) * (g.Y >> 3);
    tmp[gid] = dense[src_tile * 16u + wd];
}

// ---- B: per-brick occupancy, bounds, count --------------------------------------------------
// Word wd of a tile holds voxels z = wd>>1, y = 4*(wd&1)..+3, one byte per y row, bit = x.
__global__ void __launch_bounds__(256) analyse_bricks(const uint32_t* __restrict__ tmp, uint2* __restrict__ records,
                                                      uint32_t* __restrict__ top, unsigned long long* __restrict__ solid,
                                                      BuildGeom g, uint32_t nbricks) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint32_t brick = (uint32_t)(((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5); // 64-bit: nbricks * 32 passes 2^32 at 2^27 bricks
    if (brick >= nbricks) return;
    const uint32_t words = g.tpb * 16u;
    int mnx = 255, mny = 255, mnz = 255, mxx = -1, mxy = -1, mxz = -1;
    uint32_t count = 0;
    for (uint32_t wi = lane; wi < words; wi += 32u) {
        const uint32_t v = tmp[(uint64_t)brick * words + wi];
        if (!v) continue;
        count += __popc(v);
        const uint32_t sub = wi >> 4, wd = wi & 15u;
        const int ox = (int)((sub % g.f8) * 8u), oy = (int)(((sub / g.f8) % g.f8) * 8u + ((wd & 1u) << 2)),
                  oz = (int)((sub / (g.f8 * g.f8)) * 8u + (wd >> 1));
        const uint32_t xm = (v | (v >> 8) | (v >> 16) | (v >> 24)) & 255u;
        const uint32_t ym = ((v & 0xFFu) ? 1u : 0u) | ((v & 0xFF00u) ? 2u : 0u) | ((v & 0xFF0000u) ? 4u : 0u) | ((v & 0xFF000000u) ? 8u : 0u);
        mnx = min(mnx, ox + (__ffs(xm) - 1)); mxx = max(mxx, ox + (31 - __clz(xm)));
        mny = min(mny, oy + (__ffs(ym) - 1)); mxy = max(mxy, oy + (31 - __clz(ym)));
        mnz = min(mnz, oz); mxz = max(mxz, oz);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        mnx = min(mnx, __shfl_xor_sync(0xFFFFFFFFu, mnx, o)); mxx = max(mxx, __shfl_xor_sync(0xFFFFFFFFu, mxx, o));
        mny = min(mny, __shfl_xor_sync(0xFFFFFFFFu, mny, o)); mxy = max(mxy, __shfl_xor_sync(0xFFFFFFFFu, mxy, o));
        mnz = min(mnz, __shfl_xor_sync(0xFFFFFFFFu, mnz, o)); mxz = max(mxz, __shfl_xor_sync(0xFFFFFFFFu, mxz, o));
        count += __shfl_xor_sync(0xFFFFFFFFu, count, o);
    }
    if (lane == 0) {
        uint32_t packed = 0;
        if (count) {
            packed = (uint32_t)mnx | ((uint32_t)mny << 5) | ((uint32_t)mnz << 10) | ((uint32_t)mxx << 15) | ((uint32_t)mxy << 20) |
                     ((uint32_t)mxz << 25);
            atomicOr(&top[brick >> 5], 1u << (brick & 31u));
            atomicAdd(solid, (unsigned long long)count);
        }
        records[brick] = make_uint2(packed, 0xFFFFFFFFu);
    }
}

// ---- C: exclusive scan of the occupancy bits (popcount per word) ----------------------------
constexpr int kScanBlock = 1024; // words per block
__global__ void __launch_bounds__(kScanBlock) scan_block_sums(const uint32_t* __restrict__ top, uint32_t nwords, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t warp_sums[32];
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
    uint32_t v = i < nwords ? __popc(top[i]) : 0u;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xFFFFFFFFu, v, o);
    if ((threadIdx.x & 31) == 0) warp_sums[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
        uint32_t s = warp_sums[threadIdx.x];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, o);
        if (threadIdx.x == 0) block_sums[blockIdx.x] = s;
    }
}
__global__ void scan_block_offsets(uint32_t* __restrict__ block_sums, uint32_t nblocks, unsigned long long* __restrict__ total) {
    // single thread block; nblocks is small (2048 at 8k x 512 x 8k)
    __shared__ uint32_t carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (uint32_t base = 0; base < nblocks; base += blockDim.x) {
        const uint32_t i = base + threadIdx.x;
        const uint32_t v = i < nblocks ? block_sums[i] : 0u;
        // Hillis-Steele inclusive scan in shared memory
        __shared__ uint32_t buf[1024];
        buf[threadIdx.x] = v;
        __syncthreads();
        for (uint32_t o = 1; o < blockDim.x; o <<= 1) {
            const uint32_t add = threadIdx.x >= o ? buf[threadIdx.x - o] : 0u;
            __syncthreads();
            buf[threadIdx.x] += add;
            __syncthreads();
        }
        if (i < nblocks) block_sums[i] = carry + buf[threadIdx.x] - v;
        __syncthreads();
        if (threadIdx.x == blockDim.x - 1) carry += buf[threadIdx.x];
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}
__global__ void __launch_bounds__(kScanBlock) scan_assign_slots(const uint32_t* __restrict__ top, uint32_t nwords,
                                                                const uint32_t* __restrict__ block_offsets, uint2* __restrict__ records) {
    __shared__ uint32_t warp_sums[32];
    const uint32_t i = blockIdx.x * kScanBlock + threadIdx.x;
    const uint32_t word = i < nwords ? top[i] : 0u;
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t incl = __popc(word);
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, incl, o);
        if (lane >= (uint32_t)o) incl += n;
    }
    if (lane == 31u) warp_sums[warp] = incl;
    __syncthreads();
    if (warp == 0) {
        uint32_t s = warp_sums[lane];
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t n = __shfl_up_sync(0xFFFFFFFFu, s, o);
            if (lane >= (uint32_t)o) s += n;
        }
        warp_sums[lane] = s; // inclusive over warps
    }
    __syncthreads();
    uint32_t base = block_offsets[blockIdx.x] + (warp ? warp_sums[warp - 1] : 0u) + incl - __popc(word);
    uint32_t m = word;
    while (m) {
        const uint32_t b = __ffs(m) - 1;
        m &= m - 1;
        records[(uint64_t)i * 32u + b].y = base++;
    }
}

// records from packed bounds alone (world assembled from the parts of a partitioned build); slots follow from the scan
__global__ void __launch_bounds__(256) records_from_bounds(const uint32_t* __restrict__ bounds, uint2* __restrict__ records, uint32_t nbricks) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nbricks) records[i] = make_uint2(bounds[i], 0xFFFFFFFFu);
}
__global__ void __launch_bounds__(256) bounds_from_records(const uint2* __restrict__ records, uint32_t* __restrict__ bounds, uint32_t nbricks) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nbricks) bounds[i] = records[i].x;
}

// ---- D: compaction ---------------------------------------------------------------------------
__global__ void __launch_bounds__(256) compact_bricks(const uint32_t* __restrict__ tmp, const uint2* __restrict__ records,
                                                      uint32_t* __restrict__ pool, uint32_t words, uint64_t total_words) {
    // one thread per 16-byte chunk
    const uint64_t gid = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid * 4u >= total_words) return;
    const uint32_t brick = (uint32_t)((gid * 4u) / words);
    const uint32_t off = (uint32_t)((gid * 4u) % words);
    const uint32_t slot = records[brick].y;
    if (slot == 0xFFFFFFFFu) return;
    const uint4 v = *reinterpret_cast<const uint4*>(tmp + (uint64_t)brick * words + off);
    *reinterpret_cast<uint4*>(pool + (uint64_t)slot * words + off) = v;
}

// ---- E: traversal layout: padded x-fastest linear grid (border bits = 1) ----------------------
__global__ void __launch_bounds__(256) make_padded(const uint32_t* __restrict__ top, const uint2* __restrict__ records,
                                                   uint32_t* __restrict__ ptop, uint2* __restrict__ precords, BuildGeom g, uint64_t pcells) {
    const uint64_t wi = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; // one thread per 32-bit word of ptop
    if (wi * 32u >= pcells) return;
    const uint32_t pnx = g.cx + 2u, pny = g.cy + 2u, pnz = g.cz + 2u;
    const uint32_t tW = g.cx >> 3, tH = g.cy >> 3;
    uint32_t word = 0;
    for (uint32_t b = 0; b < 32u; ++b) {
        const uint64_t p = wi * 32u + b;
        if (p >= pcells) break;
        const uint32_t px = (uint32_t)(p % pnx), py = (uint32_t)((p / pnx) % pny), pz = (uint32_t)(p / ((uint64_t)pnx * pny));
        uint2 rec = make_uint2(0u, 0xFFFFFFFFu);
        bool bit = true; // border
        if (px >= 1u && px <= g.cx && py >= 1u && py <= g.cy && pz >= 1u && pz <= g.cz && pz < pnz) {
            const uint32_t x = px - 1u, y = py - 1u, z = pz - 1u;
            const uint32_t t = ((x >> 3) + (y >> 3) * tW + (z >> 3) * tW * tH) * 512u + (x & 7u) + ((y & 7u) << 3) + ((z & 7u) << 6);
            bit = (top[t >> 5] >> (t & 31u)) & 1u;
            rec = records[t];
        }
        word |= (uint32_t)bit << b;
        precords[p] = rec;
    }
    ptop[wi] = word;
}

// ---- export (parity) -------------------------------------------------------------------------
__global__ void export_bricks(const uint2* __restrict__ precords, const uint32_t* __restrict__ top, const uint32_t* __restrict__ pool,
                              uint32_t words, uint32_t nbricks, BuildGeom g, uint32_t* __restrict__ brick_words, float* __restrict__ bounds6) {
    const uint32_t brick = blockIdx.x;
    if (brick >= nbricks) return;
    const bool occ = (top[brick >> 5] >> (brick & 31u)) & 1u;
    uint32_t bx, by, bz;
    brick_coords(brick, g, bx, by, bz);
    const uint2 rec = precords[((uint64_t)(bz + 1u) * (g.cy + 2u) + (by + 1u)) * (g.cx + 2u) + (bx + 1u)];
    if (brick_words)
        for (uint32_t i = threadIdx.x; i < words; i += blockDim.x)
            brick_words[(uint64_t)brick * words + i] = occ ? pool[(uint64_t)rec.y * words + i] : 0u;
    if (bounds6 && threadIdx.x == 0) {
        float* o = bounds6 + (uint64_t)brick * 6u;
        if (occ) {
            const uint32_t b = rec.x;
            o[0] = (float)(b & 31u); o[1] = (float)((b >> 5) & 31u); o[2] = (float)((b >> 10) & 31u);
            o[3] = (float)((b >> 15) & 31u); o[4] = (float)((b >> 20) & 31u); o[5] = (float)((b >> 25) & 31u);
        } else {
            o[0] = o[1] = o[2] = 0.0f; o[3] = o[4] = o[5] = -1.0f;
        }
    }
}

// the same for a list of bricks given by coarse coordinates (parity spot checks on worlds too large to export whole)
__global__ void export_selected(const uint2* __restrict__ precords, const uint32_t* __restrict__ ptop, const uint32_t* __restrict__ top,
                                const uint32_t* __restrict__ pool, uint32_t words, const uint32_t* __restrict__ coords, uint32_t n, BuildGeom g,
                                uint8_t* __restrict__ occupied, uint32_t* __restrict__ brick_words, float* __restrict__ bounds6) {
    const uint32_t i = blockIdx.x;
    if (i >= n) return;
    const uint32_t bx = coords[3 * i], by = coords[3 * i + 1], bz = coords[3 * i + 2];
    const uint64_t p = ((uint64_t)(bz + 1u) * (g.cy + 2u) + (by + 1u)) * (g.cx + 2u) + (bx + 1u);
    const bool occ = (ptop[p >> 5] >> (p & 31u)) & 1u;
    const uint2 rec = precords[p];
    if (occupied && threadIdx.x == 0) {
        // both copies of the top-level bit: the traversal grid's and the reference-indexed one (bit 1)
        const uint64_t t = (uint64_t)((bx >> 3) + (by >> 3) * (g.cx >> 3) + (bz >> 3) * (g.cx >> 3) * (g.cy >> 3)) * 512u + (bx & 7u) + ((by & 7u) << 3) + ((bz & 7u) << 6);
        occupied[i] = (uint8_t)((occ ? 1u : 0u) | (((top[t >> 5] >> (t & 31u)) & 1u) << 1));
    }
    if (brick_words)
        for (uint32_t k = threadIdx.x; k < words; k += blockDim.x)
            brick_words[(uint64_t)i * words + k] = occ ? pool[(uint64_t)rec.y * words + k] : 0u;
    if (bounds6 && threadIdx.x == 0) {
        float* o = bounds6 + (uint64_t)i * 6u;
        if (occ) {
            const uint32_t b = rec.x;
            o[0] = (float)(b & 31u); o[1] = (float)((b >> 5) & 31u); o[2] = (float)((b >> 10) & 31u);
            o[3] = (float)((b >> 15) & 31u); o[4] = (float)((b >> 20) & 31u); o[5] = (float)((b >> 25) & 31u);
        } else {
            o[0] = o[1] = o[2] = 0.0f; o[3] = o[4] = o[5] = -1.0f;
        }
    }
}

} // namespace vx

// =============================================================================================
// host side
// =============================================================================================
namespace vx {

static thread_local std::string g_last_error;
void set_error(const std::string& msg) { g_last_error = msg; }
int cuda_fail(cudaError_t e, const char* what, const char* file, int line) {
    char buf[512];
    snprintf(buf, sizeof buf, "CUDA error %d (%s) in %s at %s:%d", (int)e, cudaGetErrorString(e), what, file, line);
    g_last_error = buf;
    cudaGetLastError(); // clear the sticky-less error state
    return VX_ERR_CUDA;
}

static int check_dims(uint32_t X, uint32_t Y, uint32_t Z, int factor) {
    VX_REQUIRE(factor >= 8 && factor <= 32 && factor % 8 == 0, "factor must be 8, 16, 24 or 32");
    VX_REQUIRE(X && Y && Z && X % factor == 0 && Y % factor == 0 && Z % factor == 0, "X, Y, Z must be non-zero multiples of factor");
    VX_REQUIRE((X / factor) % 8 == 0 && (Y / factor) % 8 == 0 && (Z / factor) % 8 == 0,
               "X/factor, Y/factor, Z/factor must be multiples of 8 (reference tiled coarse index)");
    VX_REQUIRE(X / factor <= 65535 && Y / factor <= 65535 && Z / factor <= 65535, "coarse dimensions exceed uint16 (reference VoxelBuffer)");
    VX_REQUIRE((uint64_t)(X / factor) * (Y / factor) * (Z / factor) < (1ull << 32) / 1, "too many bricks for a 32-bit brick index");
    return VX_OK;
}

// Stages B-D on `d_tmp` (w->nbricks bricks, brick-major): top-level bits, records (bounds + pool slot), compact pool, counters.
static int build_bricks(vx_world* w, const uint32_t* d_tmp, const BuildGeom& g) {
    const uint32_t nb = (uint32_t)w->nbricks;
    const uint32_t words = w->words_per_brick;
    const uint32_t nwords_top = (nb + 31) / 32;
    unsigned long long* d_counters = nullptr; // [0] solid voxels, [1] occupied bricks
    uint32_t* d_block_sums = nullptr;
    VX_CUDA(cudaMalloc(&w->d_top, (size_t)nwords_top * 4));
    VX_CUDA(cudaMemset(w->d_top, 0, (size_t)nwords_top * 4));
    VX_CUDA(cudaMalloc(&w->d_records, (size_t)nb * sizeof(uint2)));
    VX_CUDA(cudaMalloc(&d_counters, 16));
    VX_CUDA(cudaMemset(d_counters, 0, 16));
    analyse_bricks<<<(uint32_t)(((uint64_t)nb * 32 + 255) / 256), 256>>>(d_tmp, w->d_records, w->d_top, d_counters, g, nb);
    VX_CUDA(cudaGetLastError());
    const uint32_t nblocks = (nwords_top + kScanBlock - 1) / kScanBlock;
    VX_CUDA(cudaMalloc(&d_block_sums, (size_t)nblocks * 4));
    scan_block_sums<<<nblocks, kScanBlock>>>(w->d_top, nwords_top, d_block_sums);
    scan_block_offsets<<<1, 1024>>>(d_block_sums, nblocks, d_counters + 1);
    scan_assign_slots<<<nblocks, kScanBlock>>>(w->d_top, nwords_top, d_block_sums, w->d_records);
    VX_CUDA(cudaGetLastError());
    unsigned long long h[2] = {0, 0};
    VX_CUDA(cudaMemcpy(h, d_counters, 16, cudaMemcpyDeviceToHost));
    w->solid_voxels = h[0];
    w->occupied = h[1];
    const size_t pool_words = (size_t)(w->occupied ? w->occupied : 1) * words;
    VX_CUDA(cudaMalloc(&w->d_pool, pool_words * 4));
    w->pool_capacity = w->occupied ? w->occupied : 1; w->pool_used = w->occupied;
    if (!w->occupied) VX_CUDA(cudaMemset(w->d_pool, 0, pool_words * 4));
    const uint64_t total_words = (uint64_t)nb * words;
    compact_bricks<<<(uint32_t)((total_words / 4 + 255) / 256), 256>>>(d_tmp, w->d_records, w->d_pool, words, total_words);
    VX_CUDA(cudaGetLastError());
    cudaFree(d_counters);
    cudaFree(d_block_sums);
    return VX_OK;
}

// Stage E: the traversal copies (padded linear grid) from top bits + records; the tiled-order records are freed.
static int build_traversal(vx_world* w, const BuildGeom& g) {
    const uint64_t pcells = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
    VX_REQUIRE(pcells < (1ull << 31), "padded coarse grid exceeds a 31-bit cell index");
    VX_CUDA(cudaMalloc(&w->d_ptop, (size_t)((pcells + 31) / 32) * 4));
    VX_CUDA(cudaMalloc(&w->d_precords, (size_t)pcells * sizeof(uint2)));
    make_padded<<<(uint32_t)(((pcells + 31) / 32 + 255) / 256), 256>>>(w->d_top, w->d_records, w->d_ptop, w->d_precords, g, pcells);
    VX_CUDA(cudaGetLastError());
    VX_CUDA(cudaDeviceSynchronize());
    cudaFree(w->d_records); w->d_records = nullptr;
    return VX_OK;
}

static int finish_build(vx_world* w, const uint32_t* d_tmp, const BuildGeom& g) {
    int rc = build_bricks(w, d_tmp, g);
    if (rc) return rc;
    return build_traversal(w, g);
}

static void free_world(vx_world* w) {
    if (!w) return;
    cudaFree(w->d_top); cudaFree(w->d_records); cudaFree(w->d_pool); cudaFree(w->d_ptop); cudaFree(w->d_precords);
    delete w;
}

static int build_common(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, uint32_t x0, uint32_t y0, uint32_t z0,
                        int device, const uint32_t* host_dense, uint32_t flags) {
    VX_REQUIRE(out != nullptr, "out is NULL");
    *out = nullptr;
    int rc = check_dims(X, Y, Z, factor);
    if (rc) return rc;
    VX_REQUIRE((uint64_t)x0 + X < (1ull << 31) && (uint64_t)y0 + Y < (1ull << 31) && (uint64_t)z0 + Z < (1ull << 31), "window exceeds int range");
    DeviceGuard guard(device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    vx_world* w = new vx_world();
    w->device = device; w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    w->cx = X / factor; w->cy = Y / factor; w->cz = Z / factor;
    w->nbricks = (uint64_t)w->cx * w->cy * w->cz;
    w->words_per_brick = (uint32_t)factor * factor * factor / 32;
    BuildGeom g;
    g.cx = w->cx; g.cy = w->cy; g.cz = w->cz; g.f8 = factor / 8; g.tpb = g.f8 * g.f8 * g.f8;
    g.x0 = x0; g.y0 = y0; g.z0 = z0; g.X = X; g.Y = Y;
    const uint64_t ntiles = w->nbricks * g.tpb;
    uint32_t* d_tmp = nullptr;
    uint32_t* d_dense = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr, em = nullptr;
    auto cleanup = [&]() { cudaFree(d_tmp); cudaFree(d_dense); if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); if (em) cudaEventDestroy(em); };
#define VX_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); free_world(w); return cuda_fail(e_, #call, __FILE__, __LINE__); } } while (0)
    VX_TRY(cudaMalloc(&d_tmp, (size_t)ntiles * 64));
    VX_TRY(cudaEventCreate(&e0));
    VX_TRY(cudaEventCreate(&e1));
    VX_TRY(cudaEventCreate(&em));
    if (host_dense) {
        VX_TRY(cudaMalloc(&d_dense, (size_t)ntiles * 64));
        VX_TRY(cudaMemcpy(d_dense, host_dense, (size_t)ntiles * 64, cudaMemcpyHostToDevice));
    }
    VX_TRY(cudaEventRecord(e0));
    if (host_dense) {
        dense_tiles<<<(uint32_t)((ntiles * 16 + 255) / 256), 256>>>(d_tmp, d_dense, g, ntiles);
    } else {
        int sms = 148;
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
        const uint64_t want = (ntiles + 7) / 8; // 8 warps per block
        const uint32_t blocks = (uint32_t)(want < (uint64_t)sms * 64 ? want : (uint64_t)sms * 64);
        if (flags & 1u) noise_tiles<false><<<blocks, 256>>>(d_tmp, g, ntiles, 0);
        else noise_tiles<true><<<blocks, 256>>>(d_tmp, g, ntiles, 0);
    }
    VX_TRY(cudaGetLastError());
    VX_TRY(cudaEventRecord(em));
    rc = finish_build(w, d_tmp, g);
    if (rc) { cleanup(); free_world(w); return rc; }
    VX_TRY(cudaEventRecord(e1));
    VX_TRY(cudaEventSynchronize(e1));
    VX_TRY(cudaEventElapsedTime(&w->build_ms, e0, e1));
    VX_TRY(cudaEventElapsedTime(&w->tiles_ms, e0, em));
#undef VX_TRY
    cleanup();
    *out = w;
    return VX_OK;
}

} // namespace vx

using namespace vx;

extern "C" const char* vx_last_error(void) { return g_last_error.c_str(); }
extern "C" int vx_version(void) { return VXRT_VERSION; }
extern "C" int vx_device_count(int* count) {
    VX_REQUIRE(count != nullptr, "count is NULL");
    *count = 0;
    VX_CUDA(cudaGetDeviceCount(count));
    return VX_OK;
}

extern "C" int vx_world_build_noise(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, uint32_t x0, uint32_t y0,
                                    uint32_t z0, int device) {
    return build_common(out, X, Y, Z, factor, x0, y0, z0, device, nullptr, 0u);
}
// Validation build: every voxel runs all 32 octaves (no early-out).  Not in vxrt.h's main surface;
// declared there under "diagnostics".
extern "C" int vx_world_build_noise_full(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, uint32_t x0, uint32_t y0,
                                         uint32_t z0, int device) {
    return build_common(out, X, Y, Z, factor, x0, y0, z0, device, nullptr, 1u);
}
extern "C" int vx_world_from_dense_bits(vx_world** out, const uint32_t* host_tiled_bits, uint32_t X, uint32_t Y, uint32_t Z, int factor,
                                        int device) {
    VX_REQUIRE(host_tiled_bits != nullptr, "host_tiled_bits is NULL");
    return build_common(out, X, Y, Z, factor, 0, 0, 0, device, host_tiled_bits, 0u);
}
extern "C" int vx_world_destroy(vx_world* w) {
    if (!w) return VX_OK;
    DeviceGuard guard(w->device);
    free_world(w);
    return VX_OK;
}
extern "C" int vx_world_get_info(const vx_world* w, vx_world_info* info) {
    VX_REQUIRE(w && info, "NULL argument");
    info->dims[0] = w->X; info->dims[1] = w->Y; info->dims[2] = w->Z;
    info->coarse[0] = w->cx; info->coarse[1] = w->cy; info->coarse[2] = w->cz;
    info->factor = w->factor; info->device = w->device;
    info->bricks = w->nbricks; info->occupied_bricks = w->occupied; info->solid_voxels = w->solid_voxels;
    info->bytes_top = (w->nbricks + 31) / 32 * 4 + ((uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2) + 31) / 32 * 4; info->bytes_records = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2) * 8;
    info->bytes_pool = w->occupied * w->words_per_brick * 4;
    info->build_ms = w->build_ms;
    info->tiles_ms = w->tiles_ms;
    return VX_OK;
}
extern "C" int vx_world_download(const vx_world* w, uint32_t* top_words, uint32_t* brick_words, float* bounds6) {
    VX_REQUIRE(w != nullptr, "world is NULL");
    DeviceGuard guard(w->device);
    const uint32_t nb = (uint32_t)w->nbricks;
    if (top_words) VX_CUDA(cudaMemcpy(top_words, w->d_top, (size_t)((nb + 31) / 32) * 4, cudaMemcpyDeviceToHost));
    if (brick_words || bounds6) {
        uint32_t* d_bw = nullptr; float* d_b6 = nullptr;
        if (brick_words) VX_CUDA(cudaMalloc(&d_bw, (size_t)nb * w->words_per_brick * 4));
        if (bounds6) VX_CUDA(cudaMalloc(&d_b6, (size_t)nb * 24));
        BuildGeom g{};
        g.cx = w->cx; g.cy = w->cy; g.cz = w->cz; g.f8 = w->factor / 8; g.tpb = g.f8 * g.f8 * g.f8;
        export_bricks<<<nb, 32>>>(w->d_precords, w->d_top, w->d_pool, w->words_per_brick, nb, g, d_bw, d_b6);
        VX_CUDA(cudaGetLastError());
        if (brick_words) VX_CUDA(cudaMemcpy(brick_words, d_bw, (size_t)nb * w->words_per_brick * 4, cudaMemcpyDeviceToHost));
        if (bounds6) VX_CUDA(cudaMemcpy(bounds6, d_b6, (size_t)nb * 24, cudaMemcpyDeviceToHost));
        cudaFree(d_bw); cudaFree(d_b6);
    }
    return VX_OK;
}

extern "C" int vx_world_download_bricks(const vx_world* w, const uint32_t* coarse_xyz, size_t n, uint8_t* occupied, uint32_t* brick_words,
                                        float* bounds6) {
    VX_REQUIRE(w && (coarse_xyz || n == 0), "NULL argument");
    VX_REQUIRE(n < (1ull << 31), "too many bricks in one call");
    if (n == 0) return VX_OK;
    for (size_t i = 0; i < n; ++i)
        VX_REQUIRE(coarse_xyz[3 * i] < w->cx && coarse_xyz[3 * i + 1] < w->cy && coarse_xyz[3 * i + 2] < w->cz, "brick coordinates outside the coarse grid");
    DeviceGuard guard(w->device);
    uint32_t* d_xyz = nullptr; uint8_t* d_occ = nullptr; uint32_t* d_bw = nullptr; float* d_b6 = nullptr;
    auto cleanup = [&]() { cudaFree(d_xyz); cudaFree(d_occ); cudaFree(d_bw); cudaFree(d_b6); };
#define VX_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return cuda_fail(e_, #call, __FILE__, __LINE__); } } while (0)
    VX_TRY(cudaMalloc(&d_xyz, n * 12));
    VX_TRY(cudaMemcpy(d_xyz, coarse_xyz, n * 12, cudaMemcpyHostToDevice));
    if (occupied) VX_TRY(cudaMalloc(&d_occ, n));
    if (brick_words) VX_TRY(cudaMalloc(&d_bw, n * w->words_per_brick * 4));
    if (bounds6) VX_TRY(cudaMalloc(&d_b6, n * 24));
    BuildGeom g{};
    g.cx = w->cx; g.cy = w->cy; g.cz = w->cz; g.f8 = w->factor / 8; g.tpb = g.f8 * g.f8 * g.f8;
    export_selected<<<(uint32_t)n, 32>>>(w->d_precords, w->d_ptop, w->d_top, w->d_pool, w->words_per_brick, d_xyz, (uint32_t)n, g, d_occ, d_bw, d_b6);
    VX_TRY(cudaGetLastError());
    if (occupied) VX_TRY(cudaMemcpy(occupied, d_occ, n, cudaMemcpyDeviceToHost));
    if (brick_words) VX_TRY(cudaMemcpy(brick_words, d_bw, n * w->words_per_brick * 4, cudaMemcpyDeviceToHost));
    if (bounds6) VX_TRY(cudaMemcpy(bounds6, d_b6, n * 24, cudaMemcpyDeviceToHost));
#undef VX_TRY
    cleanup();
    return VX_OK;
}

// ---- partitioned build (BASELINE configs[4]) ------------------------------------------------------------------
// The coarse grid's 8-cell z-tiles are dealt to `slab_count` builders in contiguous runs; in the reference's tiled coarse
// index a run of z-tiles is a contiguous range of bricks, and pool slots follow brick order, so the slabs' top-level bits,
// bounds and compact pools simply concatenate into what the single-device build produces, word for word.  Each builder runs
// the voxel kernel and stages B-D on its slab (vx_slab_build_noise), the parts travel between the devices (the caller's job:
// torch.distributed / NCCL all-gather of three flat arrays), and vx_world_from_parts re-derives the slots by the same scan
// and builds the traversal copies.
struct vx_slab {
    vx_world* part = nullptr; // nbricks = the slab's bricks; d_top, d_records, d_pool valid; no traversal copies
    uint64_t brick_begin = 0;
    float build_ms = 0.0f;
};

static void slab_range(uint32_t cx, uint32_t cy, uint32_t cz, uint32_t index, uint32_t count, uint64_t& begin, uint64_t& n) {
    const uint64_t zt = cz / 8u, per_zt = (uint64_t)(cx / 8u) * (cy / 8u) * 512u;
    const uint64_t z0 = zt * index / count, z1 = zt * (index + 1u) / count;
    begin = z0 * per_zt; n = (z1 - z0) * per_zt;
}

extern "C" int vx_slab_build_noise(vx_slab** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, uint32_t x0, uint32_t y0, uint32_t z0,
                                   uint32_t slab_index, uint32_t slab_count, int device) {
    VX_REQUIRE(out != nullptr, "out is NULL");
    *out = nullptr;
    int rc = check_dims(X, Y, Z, factor);
    if (rc) return rc;
    VX_REQUIRE(slab_count >= 1 && slab_index < slab_count, "slab_index must be below slab_count");
    VX_REQUIRE((uint64_t)x0 + X < (1ull << 31) && (uint64_t)y0 + Y < (1ull << 31) && (uint64_t)z0 + Z < (1ull << 31), "window exceeds int range");
    DeviceGuard guard(device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    vx_world* w = new vx_world();
    w->device = device; w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    w->cx = X / factor; w->cy = Y / factor; w->cz = Z / factor;
    w->words_per_brick = (uint32_t)factor * factor * factor / 32;
    uint64_t begin = 0, n = 0;
    slab_range(w->cx, w->cy, w->cz, slab_index, slab_count, begin, n);
    w->nbricks = n;
    vx_slab* s = new vx_slab();
    s->part = w; s->brick_begin = begin;
    if (n == 0) { *out = s; return VX_OK; } // more builders than z-tiles: an empty slab
    BuildGeom g;
    g.cx = w->cx; g.cy = w->cy; g.cz = w->cz; g.f8 = factor / 8; g.tpb = g.f8 * g.f8 * g.f8;
    g.x0 = x0; g.y0 = y0; g.z0 = z0; g.X = X; g.Y = Y;
    const uint64_t ntiles = n * g.tpb;
    uint32_t* d_tmp = nullptr;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    auto fail = [&](int code) { cudaFree(d_tmp); if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); free_world(w); delete s; return code; };
#define VX_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(cuda_fail(e_, #call, __FILE__, __LINE__)); } while (0)
    VX_TRY(cudaMalloc(&d_tmp, (size_t)ntiles * 64));
    VX_TRY(cudaEventCreate(&e0));
    VX_TRY(cudaEventCreate(&e1));
    VX_TRY(cudaEventRecord(e0));
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const uint64_t want = (ntiles + 7) / 8;
    const uint32_t blocks = (uint32_t)(want < (uint64_t)sms * 64 ? want : (uint64_t)sms * 64);
    noise_tiles<true><<<blocks, 256>>>(d_tmp, g, ntiles, begin * g.tpb);
    VX_TRY(cudaGetLastError());
    rc = build_bricks(w, d_tmp, g);
    if (rc) return fail(rc);
    VX_TRY(cudaEventRecord(e1));
    VX_TRY(cudaEventSynchronize(e1));
    VX_TRY(cudaEventElapsedTime(&s->build_ms, e0, e1));
#undef VX_TRY
    cudaFree(d_tmp); cudaEventDestroy(e0); cudaEventDestroy(e1);
    *out = s;
    return VX_OK;
}

extern "C" int vx_slab_get_info(const vx_slab* s, vx_slab_info* info) {
    VX_REQUIRE(s && info, "NULL argument");
    info->brick_begin = s->brick_begin; info->brick_count = s->part->nbricks;
    info->occupied_bricks = s->part->occupied; info->solid_voxels = s->part->solid_voxels;
    info->words_per_brick = s->part->words_per_brick; info->build_ms = s->build_ms;
    return VX_OK;
}

extern "C" int vx_slab_export(const vx_slab* s, uint32_t* d_top_words, uint32_t* d_bounds, uint32_t* d_pool_words, void* stream) {
    VX_REQUIRE(s != nullptr, "slab is NULL");
    const vx_world* w = s->part;
    if (w->nbricks == 0) return VX_OK;
    VX_REQUIRE(d_top_words && d_bounds && (d_pool_words || w->occupied == 0), "NULL argument");
    DeviceGuard guard(w->device);
    cudaStream_t cs = (cudaStream_t)stream;
    const uint32_t nb = (uint32_t)w->nbricks;
    VX_CUDA(cudaMemcpyAsync(d_top_words, w->d_top, (size_t)(nb / 32) * 4, cudaMemcpyDeviceToDevice, cs));
    bounds_from_records<<<(nb + 255) / 256, 256, 0, cs>>>(w->d_records, d_bounds, nb);
    VX_CUDA(cudaGetLastError());
    if (w->occupied) VX_CUDA(cudaMemcpyAsync(d_pool_words, w->d_pool, (size_t)w->occupied * w->words_per_brick * 4, cudaMemcpyDeviceToDevice, cs));
    return VX_OK;
}

extern "C" int vx_slab_destroy(vx_slab* s) {
    if (!s) return VX_OK;
    DeviceGuard guard(s->part->device);
    free_world(s->part);
    delete s;
    return VX_OK;
}

extern "C" int vx_world_from_parts(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, const uint32_t* d_top_words,
                                   const uint32_t* d_bounds, const uint32_t* d_pool_words, uint64_t occupied_bricks, uint64_t solid_voxels, int device) {
    VX_REQUIRE(out && d_top_words && d_bounds && (d_pool_words || occupied_bricks == 0), "NULL argument");
    *out = nullptr;
    int rc = check_dims(X, Y, Z, factor);
    if (rc) return rc;
    DeviceGuard guard(device);
    if (!guard.ok) return cuda_fail(cudaGetLastError(), "cudaSetDevice", __FILE__, __LINE__);
    vx_world* w = new vx_world();
    w->device = device; w->X = X; w->Y = Y; w->Z = Z; w->factor = factor;
    w->cx = X / factor; w->cy = Y / factor; w->cz = Z / factor;
    w->nbricks = (uint64_t)w->cx * w->cy * w->cz;
    w->words_per_brick = (uint32_t)factor * factor * factor / 32;
    w->occupied = occupied_bricks; w->solid_voxels = solid_voxels;
    BuildGeom g{};
    g.cx = w->cx; g.cy = w->cy; g.cz = w->cz; g.f8 = factor / 8; g.tpb = g.f8 * g.f8 * g.f8; g.X = X; g.Y = Y;
    const uint32_t nb = (uint32_t)w->nbricks, nwords_top = nb / 32;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    uint32_t* d_block_sums = nullptr; unsigned long long* d_total = nullptr;
    auto fail = [&](int code) { cudaFree(d_block_sums); cudaFree(d_total); if (e0) cudaEventDestroy(e0); if (e1) cudaEventDestroy(e1); free_world(w); return code; };
#define VX_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail(cuda_fail(e_, #call, __FILE__, __LINE__)); } while (0)
    VX_TRY(cudaEventCreate(&e0)); VX_TRY(cudaEventCreate(&e1));
    VX_TRY(cudaEventRecord(e0));
    VX_TRY(cudaMalloc(&w->d_top, (size_t)nwords_top * 4));
    VX_TRY(cudaMemcpy(w->d_top, d_top_words, (size_t)nwords_top * 4, cudaMemcpyDeviceToDevice));
    VX_TRY(cudaMalloc(&w->d_records, (size_t)nb * sizeof(uint2)));
    records_from_bounds<<<(nb + 255) / 256, 256>>>(d_bounds, w->d_records, nb);
    const uint32_t nblocks = (nwords_top + kScanBlock - 1) / kScanBlock;
    VX_TRY(cudaMalloc(&d_block_sums, (size_t)nblocks * 4));
    VX_TRY(cudaMalloc(&d_total, 8));
    scan_block_sums<<<nblocks, kScanBlock>>>(w->d_top, nwords_top, d_block_sums);
    scan_block_offsets<<<1, 1024>>>(d_block_sums, nblocks, d_total);
    scan_assign_slots<<<nblocks, kScanBlock>>>(w->d_top, nwords_top, d_block_sums, w->d_records);
    VX_TRY(cudaGetLastError());
    unsigned long long total = 0;
    VX_TRY(cudaMemcpy(&total, d_total, 8, cudaMemcpyDeviceToHost));
    if (total != occupied_bricks) { set_error("vx_world_from_parts: the top-level bits name a different number of bricks than the pool holds"); return fail(VX_ERR_INVALID); }
    const size_t pool_words = (size_t)(w->occupied ? w->occupied : 1) * w->words_per_brick;
    VX_TRY(cudaMalloc(&w->d_pool, pool_words * 4));
    w->pool_capacity = w->occupied ? w->occupied : 1; w->pool_used = w->occupied;
    if (w->occupied) VX_TRY(cudaMemcpy(w->d_pool, d_pool_words, pool_words * 4, cudaMemcpyDeviceToDevice));
    else VX_TRY(cudaMemset(w->d_pool, 0, pool_words * 4));
    rc = build_traversal(w, g);
    if (rc) return fail(rc);
    VX_TRY(cudaEventRecord(e1));
    VX_TRY(cudaEventSynchronize(e1));
    VX_TRY(cudaEventElapsedTime(&w->build_ms, e0, e1));
#undef VX_TRY
    cudaFree(d_block_sums); cudaFree(d_total); cudaEventDestroy(e0); cudaEventDestroy(e1);
    *out = w;
    return VX_OK;
}
