// Editing a built brickmap in place, and its on-disk form (the reference README's "modifiable
// terrain" / "chunk streaming" TODO items; the reference itself can only flip bits of a dense grid through
// BitRef::operator=, VolumeRaytracer.cu:19-36, and rebuild everything with GenerateLowresVoxelBuffer).  After an edit the
// world is, bit for bit, what a rebuild from the edited dense grid would give (top-level bits, brick masks, bounds), so
// every parity statement about traversal carries over.
#include "vx_host.h"
#include <stddef.h>

#include <cstring>
#include <vector>

using namespace vx;

namespace vx {

constexpr uint32_t kNoSlot = 0xFFFFFFFFu, kClaimed = 0xFFFFFFFEu;

struct EditGeom {
    uint32_t X, Y, Z, factor, f8, words;
    uint32_t cx, cy, cz;
};
__device__ __forceinline__ uint64_t padded_cell(const EditGeom& g, uint32_t bx, uint32_t by, uint32_t bz) {
    return ((uint64_t)(bz + 1u) * (g.cy + 2u) + (by + 1u)) * (g.cx + 2u) + (bx + 1u);
}
__device__ __forceinline__ uint32_t tiled_brick(const EditGeom& g, uint32_t bx, uint32_t by, uint32_t bz) { // GetSampleIndex, VolumeRaytracer.cuh:111-131
    return ((bx >> 3) + (by >> 3) * (g.cx >> 3) + (bz >> 3) * (g.cx >> 3) * (g.cy >> 3)) * 512u + (bx & 7u) + ((by & 7u) << 3) + ((bz & 7u) << 6);
}

// pass 1 (solid edits only): every empty brick that is about to receive a voxel gets a pool slot, once
__global__ void __launch_bounds__(256) edit_claim(const uint32_t* __restrict__ xyz, size_t n, EditGeom g, uint2* __restrict__ precords,
                                                  unsigned long long* __restrict__ counters) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (x >= g.X || y >= g.Y || z >= g.Z) return;
    uint32_t* slot = &precords[padded_cell(g, x / g.factor, y / g.factor, z / g.factor)].y;
    if (*slot == kNoSlot && atomicCAS(slot, kNoSlot, kClaimed) == kNoSlot)
        *slot = (uint32_t)atomicAdd(&counters[0], 1ull); // slots past pool_used are zero-filled already
}

// pass 2: flip the bits; remember each touched brick once
__global__ void __launch_bounds__(256) edit_apply(const uint32_t* __restrict__ xyz, size_t n, int solid, EditGeom g, const uint2* __restrict__ precords,
                                                  uint32_t* __restrict__ pool, uint32_t* __restrict__ dirty_bits, uint32_t* __restrict__ dirty_list,
                                                  unsigned long long* __restrict__ counters) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t x = xyz[3 * i], y = xyz[3 * i + 1], z = xyz[3 * i + 2];
    if (x >= g.X || y >= g.Y || z >= g.Z) return;
    const uint32_t bx = x / g.factor, by = y / g.factor, bz = z / g.factor;
    const uint64_t cell = padded_cell(g, bx, by, bz);
    const uint32_t slot = precords[cell].y;
    if (slot == kNoSlot) return; // clearing a voxel of an empty brick
    const uint32_t lx = x - bx * g.factor, ly = y - by * g.factor, lz = z - bz * g.factor;
    const uint32_t sub = (lx >> 3) + (ly >> 3) * g.f8 + (lz >> 3) * g.f8 * g.f8;
    const uint32_t bit = (lx & 7u) + ((ly & 7u) << 3) + ((lz & 7u) << 6);
    uint32_t* word = pool + (uint64_t)slot * g.words + sub * 16u + (bit >> 5);
    const uint32_t m = 1u << (bit & 31u);
    const uint32_t old = solid ? atomicOr(word, m) : atomicAnd(word, ~m);
    if (((old & m) != 0u) == (solid != 0)) return; // no change
    atomicAdd(&counters[1], solid ? 1ull : ~0ull);      // solid voxels, two's complement for -1
    if (!(atomicOr(&dirty_bits[cell >> 5], 1u << (cell & 31u)) & (1u << (cell & 31u))))
        dirty_list[atomicAdd(&counters[2], 1ull)] = (uint32_t)cell;
}

// pass 3: one warp per touched brick: bounds and occupancy from its mask (what analyse_bricks does at build time)
__global__ void __launch_bounds__(256) edit_refresh(EditGeom g, uint2* __restrict__ precords, const uint32_t* __restrict__ pool, uint32_t* __restrict__ top,
                                                    uint32_t* __restrict__ ptop, uint32_t* __restrict__ dirty_bits, const uint32_t* __restrict__ dirty_list,
                                                    unsigned long long* __restrict__ counters) {
    const uint32_t lane = threadIdx.x & 31u;
    const unsigned long long k = ((unsigned long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (k >= counters[2]) return;
    const uint32_t cell = dirty_list[k];
    const uint32_t pnx = g.cx + 2u, pny = g.cy + 2u;
    const uint32_t bx = cell % pnx - 1u, by = (cell / pnx) % pny - 1u, bz = cell / (pnx * pny) - 1u;
    const uint32_t slot = precords[cell].y;
    int mnx = 255, mny = 255, mnz = 255, mxx = -1, mxy = -1, mxz = -1;
    uint32_t count = 0;
    for (uint32_t wi = lane; wi < g.words; wi += 32u) {
        const uint32_t v = pool[(uint64_t)slot * g.words + wi];
        if (!v) continue;
        count += __popc(v);
        const uint32_t sub = wi >> 4, wd = wi & 15u;
        const int ox = (int)((sub % g.f8) * 8u), oy = (int)(((sub / g.f8) % g.f8) * 8u + ((wd & 1u) << 2)), oz = (int)((sub / (g.f8 * g.f8)) * 8u + (wd >> 1));
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
    if (lane) return;
    const uint32_t t = tiled_brick(g, bx, by, bz);
    const bool was = (top[t >> 5] >> (t & 31u)) & 1u;
    if (count) {
        precords[cell].x = (uint32_t)mnx | ((uint32_t)mny << 5) | ((uint32_t)mnz << 10) | ((uint32_t)mxx << 15) | ((uint32_t)mxy << 20) | ((uint32_t)mxz << 25);
        if (!was) { atomicOr(&top[t >> 5], 1u << (t & 31u)); atomicOr(&ptop[cell >> 5], 1u << (cell & 31u)); atomicAdd(&counters[3], 1ull); }
    } else {
        precords[cell].x = 0u; // an emptied brick keeps its (all-zero) slot for the next edit
        if (was) { atomicAnd(&top[t >> 5], ~(1u << (t & 31u))); atomicAnd(&ptop[cell >> 5], ~(1u << (cell & 31u))); atomicAdd(&counters[3], ~0ull); }
    }
    atomicAnd(&dirty_bits[cell >> 5], ~(1u << (cell & 31u)));
}

} // namespace vx

static EditGeom geom_of(const vx_world* w) {
    EditGeom g;
    g.X = w->X; g.Y = w->Y; g.Z = w->Z; g.factor = (uint32_t)w->factor; g.f8 = g.factor / 8u; g.words = w->words_per_brick;
    g.cx = w->cx; g.cy = w->cy; g.cz = w->cz;
    return g;
}

static int grow_pool(vx_world* w, uint64_t need) {
    if (need <= w->pool_capacity) return VX_OK;
    uint64_t cap = w->pool_capacity + w->pool_capacity / 2 + 64;
    if (cap < need) cap = need;
    if (cap > w->nbricks) cap = w->nbricks;
    uint32_t* p = nullptr;
    const size_t wb = (size_t)w->words_per_brick * 4;
    VX_CUDA(cudaMalloc(&p, cap * wb));
    cudaError_t e = cudaMemcpy(p, w->d_pool, w->pool_used * wb, cudaMemcpyDeviceToDevice);
    if (e == cudaSuccess) e = cudaMemset((char*)p + w->pool_used * wb, 0, (cap - w->pool_used) * wb);
    if (e != cudaSuccess) { cudaFree(p); VX_CUDA(e); }
    cudaFree(w->d_pool);
    w->d_pool = p; w->pool_capacity = cap;
    return VX_OK;
}

extern "C" int vx_world_edit(vx_world* w, const uint32_t* voxels_xyz, size_t n, int solid) {
    VX_REQUIRE(w != nullptr, "world is NULL");
    if (!n) return VX_OK;
    VX_REQUIRE(voxels_xyz != nullptr, "voxels_xyz is NULL");
    VX_REQUIRE(n < (1ull << 32), "too many edits for one call");
    DeviceGuard guard(w->device);
    VX_CUDA(cudaDeviceSynchronize()); // frames in flight read the structures this call rewrites (and may reallocate)
    if (solid) {
        const uint64_t room = w->nbricks - w->pool_used;
        int rc = grow_pool(w, w->pool_used + (n < room ? n : room)); // at most one new brick per edit
        if (rc) return rc;
    }
    const EditGeom g = geom_of(w);
    const uint64_t pcells = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
    uint32_t *d_xyz = nullptr, *d_dirty_bits = nullptr, *d_dirty_list = nullptr;
    unsigned long long* d_counters = nullptr; // [0] pool_used, [1] solid voxels delta, [2] touched bricks, [3] occupied delta
    auto cleanup = [&]() { cudaFree(d_xyz); cudaFree(d_dirty_bits); cudaFree(d_dirty_list); cudaFree(d_counters); };
#define VX_TRY(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { cleanup(); return cuda_fail(e_, #call, __FILE__, __LINE__); } } while (0)
    VX_TRY(cudaMalloc(&d_xyz, n * 12));
    VX_TRY(cudaMalloc(&d_dirty_bits, (size_t)((pcells + 31) / 32) * 4));
    VX_TRY(cudaMalloc(&d_dirty_list, n * 4));
    VX_TRY(cudaMalloc(&d_counters, 32));
    VX_TRY(cudaMemcpy(d_xyz, voxels_xyz, n * 12, cudaMemcpyHostToDevice));
    VX_TRY(cudaMemset(d_dirty_bits, 0, (size_t)((pcells + 31) / 32) * 4));
    unsigned long long h[4] = {w->pool_used, 0, 0, 0};
    VX_TRY(cudaMemcpy(d_counters, h, 32, cudaMemcpyHostToDevice));
    const unsigned blocks = (unsigned)((n + 255) / 256);
    if (solid) edit_claim<<<blocks, 256>>>(d_xyz, n, g, w->d_precords, d_counters);
    edit_apply<<<blocks, 256>>>(d_xyz, n, solid, g, w->d_precords, w->d_pool, d_dirty_bits, d_dirty_list, d_counters);
    edit_refresh<<<(unsigned)((n * 32 + 255) / 256), 256>>>(g, w->d_precords, w->d_pool, w->d_top, w->d_ptop, d_dirty_bits, d_dirty_list, d_counters);
    VX_TRY(cudaGetLastError());
    VX_TRY(cudaMemcpy(h, d_counters, 32, cudaMemcpyDeviceToHost));
#undef VX_TRY
    cleanup();
    w->pool_used = h[0];
    w->solid_voxels += h[1]; // deltas are two's complement
    w->occupied += h[3];
    return VX_OK;
}

// ---- on-disk form -----------------------------------------------------------------------------------------------------
// One file: header, then the four device arrays as they are used by traversal (so loading is four reads and four copies):
// top-level bits (tiled order), padded top-level bits, padded brick records, brick mask pool (pool_used slots).
namespace {
struct FileHeader {
    char magic[8];
    uint32_t version, X, Y, Z, factor, words_per_brick, reserved;
    uint64_t nbricks, occupied, solid_voxels, pool_used;
    uint64_t bytes_top, bytes_ptop, bytes_precords, bytes_pool;
    uint64_t checksum; // FNV-1a over the header up to this field, then the four payload sections
};
const char kMagic[8] = {'V', 'X', 'B', 'R', 'I', 'C', 'K', '1'};

uint64_t fnv1a(uint64_t h, const void* p, size_t n) {
    const uint8_t* b = (const uint8_t*)p;
    for (size_t i = 0; i < n; ++i) { h ^= b[i]; h *= 1099511628211ull; }
    return h;
}
void section_sizes(const vx_world* w, uint64_t out[4]) {
    const uint64_t pcells = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
    out[0] = (w->nbricks + 31) / 32 * 4; out[1] = (pcells + 31) / 32 * 4; out[2] = pcells * 8; out[3] = w->pool_used * w->words_per_brick * 4;
}
constexpr size_t kChunk = 64u << 20;
constexpr uint32_t kFileVersion = 2; // 2: the checksum covers the header too

// What the march relies on without checking: every border cell of the traversal grid reads as occupied, and every
// occupied interior cell's record points into the pool.  Counts violations.
__global__ void validate_traversal_grid(const uint32_t* __restrict__ ptop, const uint2* __restrict__ precords, uint32_t cx, uint32_t cy, uint32_t cz,
                                        uint64_t pcells, uint64_t pool_used, unsigned long long* __restrict__ bad) {
    const uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= pcells) return;
    const uint32_t pnx = cx + 2u, pny = cy + 2u;
    const uint32_t x = (uint32_t)(p % pnx), y = (uint32_t)((p / pnx) % pny), z = (uint32_t)(p / ((uint64_t)pnx * pny));
    const bool border = x == 0u || y == 0u || z == 0u || x == cx + 1u || y == cy + 1u || z == cz + 1u;
    const bool bit = (ptop[p >> 5] >> (p & 31u)) & 1u;
    const bool wrong = border ? !bit : (bit && (uint64_t)precords[p].y >= pool_used);
    if (wrong) atomicAdd(bad, 1ull);
}
} // namespace

extern "C" int vx_world_save(const vx_world* w, const char* path) {
    VX_REQUIRE(w && path, "NULL argument");
    DeviceGuard guard(w->device);
    VX_CUDA(cudaDeviceSynchronize());
    FILE* f = fopen(path, "wb");
    VX_REQUIRE(f != nullptr, std::string("cannot open for writing: ") + path);
    FileHeader h{};
    memcpy(h.magic, kMagic, 8);
    h.version = kFileVersion; h.X = w->X; h.Y = w->Y; h.Z = w->Z; h.factor = (uint32_t)w->factor; h.words_per_brick = w->words_per_brick;
    h.nbricks = w->nbricks; h.occupied = w->occupied; h.solid_voxels = w->solid_voxels; h.pool_used = w->pool_used;
    uint64_t sz[4];
    section_sizes(w, sz);
    h.bytes_top = sz[0]; h.bytes_ptop = sz[1]; h.bytes_precords = sz[2]; h.bytes_pool = sz[3];
    const void* src[4] = {w->d_top, w->d_ptop, w->d_precords, w->d_pool};
    std::vector<uint8_t> buf(kChunk);
    bool ok = fwrite(&h, sizeof h, 1, f) == 1;
    uint64_t sum = fnv1a(14695981039346656037ull, &h, offsetof(FileHeader, checksum));
    for (int s = 0; s < 4 && ok; ++s)
        for (uint64_t off = 0; off < sz[s] && ok; off += kChunk) {
            const size_t m = (size_t)(sz[s] - off < kChunk ? sz[s] - off : kChunk);
            cudaError_t e = cudaMemcpy(buf.data(), (const char*)src[s] + off, m, cudaMemcpyDeviceToHost);
            if (e != cudaSuccess) { fclose(f); VX_CUDA(e); }
            sum = fnv1a(sum, buf.data(), m);
            ok = fwrite(buf.data(), 1, m, f) == m;
        }
    h.checksum = sum;
    ok = ok && fseek(f, 0, SEEK_SET) == 0 && fwrite(&h, sizeof h, 1, f) == 1;
    ok = (fclose(f) == 0) && ok;
    VX_REQUIRE(ok, std::string("write failed: ") + path);
    return VX_OK;
}

extern "C" int vx_world_load(vx_world** out, const char* path, int device) {
    VX_REQUIRE(out && path, "NULL argument");
    *out = nullptr;
    FILE* f = fopen(path, "rb");
    VX_REQUIRE(f != nullptr, std::string("cannot open: ") + path);
    FileHeader h{};
    if (fread(&h, sizeof h, 1, f) != 1 || memcmp(h.magic, kMagic, 8) != 0 || h.version != kFileVersion) { fclose(f); VX_REQUIRE(false, "not a VXBRICK1 file of this version"); }
    DeviceGuard guard(device);
    vx_world* w = new vx_world();
    w->device = device; w->X = h.X; w->Y = h.Y; w->Z = h.Z; w->factor = (int)h.factor; w->words_per_brick = h.words_per_brick;
    bool sane = h.factor >= 8 && h.factor <= 32 && h.factor % 8 == 0 && h.X && h.Y && h.Z && h.X % h.factor == 0 && h.Y % h.factor == 0 && h.Z % h.factor == 0;
    if (sane) {
        w->cx = h.X / h.factor; w->cy = h.Y / h.factor; w->cz = h.Z / h.factor;
        w->nbricks = (uint64_t)w->cx * w->cy * w->cz; w->occupied = h.occupied; w->solid_voxels = h.solid_voxels;
        w->pool_used = h.pool_used; w->pool_capacity = h.pool_used ? h.pool_used : 1;
        uint64_t sz[4];
        section_sizes(w, sz);
        // the invariants the builders enforce (check_dims, finish_build): uint16 coarse dims, 32-bit brick index, 31-bit padded cell index
        const uint64_t pcells = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
        sane = w->cx <= 65535u && w->cy <= 65535u && w->cz <= 65535u && w->nbricks < (1ull << 32) && pcells < (1ull << 31) &&
               w->cx % 8 == 0 && w->cy % 8 == 0 && w->cz % 8 == 0 && h.nbricks == w->nbricks && h.words_per_brick == h.factor * h.factor * h.factor / 32 &&
               h.pool_used <= w->nbricks && h.occupied <= h.pool_used && sz[0] == h.bytes_top && sz[1] == h.bytes_ptop && sz[2] == h.bytes_precords && sz[3] == h.bytes_pool;
    }
    if (!sane) { fclose(f); delete w; VX_REQUIRE(false, "inconsistent header"); }
    const uint64_t sz[4] = {h.bytes_top, h.bytes_ptop, h.bytes_precords, h.bytes_pool};
    void** dst[4] = {(void**)&w->d_top, (void**)&w->d_ptop, (void**)&w->d_precords, (void**)&w->d_pool};
    std::vector<uint8_t> buf(kChunk);
    uint64_t sum = fnv1a(14695981039346656037ull, &h, offsetof(FileHeader, checksum));
    const char* why = nullptr;
    for (int s = 0; s < 4 && !why; ++s) {
        const size_t alloc = (size_t)(s == 3 ? w->pool_capacity * w->words_per_brick * 4 : sz[s]);
        if (cudaMalloc(dst[s], alloc) != cudaSuccess || (s == 3 && cudaMemset(*dst[s], 0, alloc) != cudaSuccess)) { why = "out of device memory"; break; }
        for (uint64_t off = 0; off < sz[s] && !why; off += kChunk) {
            const size_t m = (size_t)(sz[s] - off < kChunk ? sz[s] - off : kChunk);
            if (fread(buf.data(), 1, m, f) != m) { why = "file is truncated"; break; }
            sum = fnv1a(sum, buf.data(), m);
            if (cudaMemcpy((char*)*dst[s] + off, buf.data(), m, cudaMemcpyHostToDevice) != cudaSuccess) why = "copy to the device failed";
        }
    }
    fclose(f);
    if (!why && sum != h.checksum) why = "checksum mismatch";
    if (!why) { // a file that passes the checksum may still come from elsewhere: check what the march takes on trust
        unsigned long long* d_bad = nullptr;
        unsigned long long bad = 0;
        const uint64_t pcells = (uint64_t)(w->cx + 2) * (w->cy + 2) * (w->cz + 2);
        if (cudaMalloc(&d_bad, 8) != cudaSuccess || cudaMemset(d_bad, 0, 8) != cudaSuccess) why = "out of device memory";
        else {
            validate_traversal_grid<<<(unsigned)((pcells + 255) / 256), 256>>>(w->d_ptop, w->d_precords, w->cx, w->cy, w->cz, pcells, w->pool_used, d_bad);
            if (cudaMemcpy(&bad, d_bad, 8, cudaMemcpyDeviceToHost) != cudaSuccess) why = "validation failed to run";
            else if (bad) why = "traversal grid is inconsistent (border bits or pool slots)";
        }
        cudaFree(d_bad);
    }
    if (why) {
        cudaGetLastError();
        cudaFree(w->d_top); cudaFree(w->d_ptop); cudaFree(w->d_precords); cudaFree(w->d_pool);
        delete w;
        VX_REQUIRE(false, std::string("vx_world_load: ") + why);
    }
    *out = w;
    return VX_OK;
}
