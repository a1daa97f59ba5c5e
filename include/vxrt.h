/*
 * vxrt.h — C ABI of the B200-native brickmap voxel ray tracer (libvxrt.so).
 *
 * This is the drop-in boundary for the reference's one hot path (JoshuaLim007/VoxelEngine,
 * VoxelRT/): noise world build -> brickmap -> two-level DDA ray march -> shading.  The reference
 * has no FFI layer; the calls below replace the C++ API that VoxelApp/main.cu uses, each entry
 * citing the reference interface it stands in for (paths relative to the reference root).
 * The header-compatible C++ shim with the reference's own names lives in include/voxelrt_compat/.
 *
 * Plain C: pointers and sizes only, no CUDA or torch types.  `stream` arguments are a
 * cudaStream_t passed as void* (NULL = the legacy default stream).  Every function returns
 * VX_OK (0) or a negative error; vx_last_error() gives the message of the calling thread's last
 * failure.  There is no CPU fallback: without a CUDA device every compute entry fails with
 * VX_ERR_CUDA.
 */
#ifndef VXRT_H
#define VXRT_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VXRT_VERSION 100

enum {
    VX_OK = 0,
    VX_ERR_INVALID = -1, /* bad argument (dims not multiples of factor, factor not a multiple of 8, ...) */
    VX_ERR_CUDA = -2,    /* CUDA runtime error, message in vx_last_error() */
    VX_ERR_CAPACITY = -3 /* more rays than the tracer was created for */
};

/* How Raytrace()'s `normalize(ray)` (helper_math.h:1325-1329) is evaluated.
 * DEVICE: the device rsqrtf — what the reference's kernels do on a GPU (default).
 * HOST:   (float)(1.0/sqrt((double)x)) — CUDA's host definition of rsqrtf
 *         (crt/math_functions.hpp), i.e. what the reference computes when its source runs on a
 *         CPU; used for bit-exact parity against the CPU oracle.
 * NONE:   directions are taken as already normalised. */
enum { VX_NORM_DEVICE = 0, VX_NORM_HOST = 1, VX_NORM_NONE = 2 };

typedef struct vx_world vx_world;   /* brickmap resident on one GPU */
typedef struct vx_tracer vx_tracer; /* per-device ray buffers bound to a world */

const char* vx_last_error(void);
int vx_version(void);
int vx_device_count(int* count);

/* ------------------------------------------------------------------------------------------
 * World build.
 * vx_world_build_noise replaces  CreateVoxels(uint3)                 VoxelWorldBuilder.cuh:12-32
 *                              + GenerateLowresVoxelBuffer(buf, f)    VolumeRaytracer.cuh:379-516
 *                              + VoxelRaytracer3D::Upload*            VolumeRaytracer.cu:527-572
 * in one device-side pass (no dense host grid, no per-brick allocation).  (x0,y0,z0) selects a
 * window of the infinite noise field (the reference always builds from 0,0,0).
 * vx_world_from_dense_bits replaces GenerateLowresVoxelBuffer + Upload* for an externally
 * supplied dense grid in the reference's tiled-linear bit layout (VolumeRaytracer.cuh:111-131).
 * X,Y,Z must be multiples of `factor`, factor a multiple of 8, and X/factor, Y/factor, Z/factor
 * multiples of 8 (the reference's coarse grid uses the same 8^3-tiled index).
 * ------------------------------------------------------------------------------------------ */
int vx_world_build_noise(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor,
                         uint32_t x0, uint32_t y0, uint32_t z0, int device);
int vx_world_from_dense_bits(vx_world** out, const uint32_t* host_tiled_bits, uint32_t X, uint32_t Y,
                             uint32_t Z, int factor, int device);
int vx_world_destroy(vx_world* w);
/* Diagnostics: same as vx_world_build_noise with the exact early-out of the noise evaluation
 * disabled (every voxel runs all 32 octaves, as the reference's PopulateVoxels does). */
int vx_world_build_noise_full(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor,
                              uint32_t x0, uint32_t y0, uint32_t z0, int device);

/* Partitioned build (BASELINE configs[4]; no reference counterpart: CreateVoxels builds on one GPU).  The coarse
 * grid's 8-cell z-tiles are dealt to slab_count builders in contiguous runs; builder slab_index runs the voxel kernel and the
 * brick analysis on its slab only.  A slab's parts — top-level bits (brick_count / 32 words), packed bounds (one uint32 per
 * brick) and compact mask pool (occupied_bricks * words_per_brick words) — are contiguous pieces of the whole world's arrays
 * in brick order: vx_slab_export copies them to DEVICE buffers of the caller, the caller moves them between the devices
 * (torch.distributed / NCCL all-gather in voxelengine_b200/partition.py), and vx_world_from_parts turns the concatenation
 * (device pointers on `device`) into a world that is word for word the one vx_world_build_noise builds. */
typedef struct vx_slab vx_slab;
typedef struct vx_slab_info {
    uint64_t brick_begin, brick_count;  /* the slab's range in the reference's tiled coarse index */
    uint64_t occupied_bricks, solid_voxels;
    uint32_t words_per_brick;
    float build_ms;
} vx_slab_info;
int vx_slab_build_noise(vx_slab** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, uint32_t x0, uint32_t y0, uint32_t z0,
                        uint32_t slab_index, uint32_t slab_count, int device);
int vx_slab_get_info(const vx_slab* s, vx_slab_info* info);
int vx_slab_export(const vx_slab* s, uint32_t* d_top_words, uint32_t* d_bounds, uint32_t* d_pool_words, void* stream);
int vx_slab_destroy(vx_slab* s);
int vx_world_from_parts(vx_world** out, uint32_t X, uint32_t Y, uint32_t Z, int factor, const uint32_t* d_top_words,
                        const uint32_t* d_bounds, const uint32_t* d_pool_words, uint64_t occupied_bricks, uint64_t solid_voxels, int device);

typedef struct vx_world_info {
    uint32_t dims[3];        /* voxels */
    uint32_t coarse[3];      /* bricks */
    int32_t factor;
    int32_t device;
    uint64_t bricks;          /* coarse cells */
    uint64_t occupied_bricks; /* bricks with at least one voxel (= slots in the pool) */
    uint64_t solid_voxels;
    uint64_t bytes_top, bytes_records, bytes_pool;
    float build_ms;           /* device time of the last build (CUDA events) */
    float tiles_ms;           /* of which: the voxel kernel (noise_tiles, the PopulateVoxels replacement; dense_tiles for a given grid) */
} vx_world_info;
int vx_world_get_info(const vx_world* w, vx_world_info* info);

/* Reference-form export for parity checks: coarse occupancy bits (tiled index, (bricks+31)/32
 * words), per-brick masks indexed by coarse tiled index (bricks * factor^3/32 words, zeros for
 * empty bricks) and per-brick bounds (bricks * 6 floats: min xyz, max xyz; 0,0,0,-1,-1,-1 when
 * empty) — exactly what GenerateLowresVoxelBuffer returns.  Any pointer may be NULL. */
int vx_world_download(const vx_world* w, uint32_t* top_words, uint32_t* brick_words, float* bounds6);
/* The same export for n bricks picked by coarse coordinates (host array of n*3 uint32): occupied[i] bit 0 = the traversal
 * grid's top-level bit, bit 1 = the reference-indexed top-level bit (they must agree); masks and bounds as above, n
 * entries each.  For parity spot checks of worlds too large to export whole (8192x512x8192 is 5.5 GB in reference form). */
int vx_world_download_bricks(const vx_world* w, const uint32_t* coarse_xyz, size_t n, uint8_t* occupied, uint32_t* brick_words,
                             float* bounds6);
/* Editing and storing a built world (the reference README's "modifiable terrain" / streaming items; the
 * reference can only flip bits of the dense grid through BitRef::operator=, VolumeRaytracer.cu:19-36, and rebuild).
 * vx_world_edit sets (solid != 0) or clears the n voxels voxels_xyz[3*i..] (host array; coordinates outside the world
 * are ignored, duplicates are fine) and refreshes the touched bricks' masks, bounds and top-level bits on the device:
 * afterwards the world is bit for bit what vx_world_from_dense_bits would build from the edited dense grid.
 * Synchronous; waits for work in flight on the device first (the brick pool may move).
 * vx_world_save / vx_world_load: the built brickmap as one file ("VXBRICK1": header, top-level bits, padded traversal
 * grid, brick records, mask pool, FNV-1a checksum); a loaded world traces and edits like the one that was saved. */
int vx_world_edit(vx_world* w, const uint32_t* voxels_xyz, size_t n, int solid);
int vx_world_save(const vx_world* w, const char* path);
int vx_world_load(vx_world** out, const char* path, int device);

/* ------------------------------------------------------------------------------------------
 * Tracer.  Replaces VoxelRaytracer3D(size_t count) / ~VoxelRaytracer3D  VolumeRaytracer.cuh:318-334
 * ------------------------------------------------------------------------------------------ */
int vx_tracer_create(vx_tracer** out, vx_world* w, size_t max_rays);
int vx_tracer_destroy(vx_tracer* t);

/* Per-ray result planes (structure of arrays).  Any pointer may be NULL.
 * hit/pos/normal/steps are what the reference's `dispatch` kernel returns
 * (VolumeRaytracer.cu:95-117: pos = +inf on a miss, normal un-negated); the rest are the locals
 * of Raytrace() the parity contract compares. */
typedef struct vx_hits {
    uint8_t* hit;          /* n */
    float* pos;            /* n*3 */
    float* normal;         /* n*3 */
    int32_t* steps;        /* n */
    int32_t* voxel;        /* n*3: coarse cell*factor + fine cell, -1 on a miss */
    uint32_t* brick;       /* n: tiled coarse index (VolumeRaytracer.cu:419), 0xFFFFFFFF on a miss */
    int32_t* coarse_steps; /* n */
    int32_t* fine_steps;   /* n */
    uint32_t* counters;    /* n*3: C coarse cells examined, K occupied cells examined, E bricks entered */
} vx_hits;

typedef struct vx_trace_opts {
    int32_t norm_mode; /* VX_NORM_* */
    int32_t max_steps; /* Raytrace(maxSteps, ...); 0 = MAX_STEPS = 2048 (VolumeRaytracer.cuh:235) */
} vx_trace_opts;

/* Batch rays, host buffers: H2D of origins/dirs, kernel, D2H of the requested planes.
 * Replaces RayTraceResults<float3> VoxelRaytracer3D::Raytrace(vector<float3>, vector<float3>)
 * (VolumeRaytracer.cu:574-618).  opts may be NULL (defaults). */
int vx_trace(vx_tracer* t, const float* origins_xyz, const float* dirs_xyz, size_t n,
             const vx_trace_opts* opts, const vx_hits* out_host);
/* Same, all pointers are device pointers on the tracer's device; asynchronous on `stream`. */
int vx_trace_device(vx_tracer* t, const float* d_origins_xyz, const float* d_dirs_xyz, size_t n,
                    const vx_trace_opts* opts, const vx_hits* out_dev, void* stream);
/* Host post-processing of the batch API (VolumeRaytracer.cu:601-614): valid / distance. */
int vx_hits_postprocess(const float* origins_xyz, const float* pos, size_t n, uint8_t* valid, float* distance);

/* ------------------------------------------------------------------------------------------
 * Screen render.  Replaces Graphics::RenderScreen (Renderer.cu:305-328) and the process globals
 * behind SetEnvironment / SetFOV (Renderer.cu:278-303), which are passed explicitly here.
 * ------------------------------------------------------------------------------------------ */
typedef struct vx_camera {
    float origin[3], fwd[3], up[3], right[3]; /* as produced by Graphics::GetDirections */
    float fov_deg;                            /* SetFOV */
    float tan_half;                           /* 0: tanf(fov/2) evaluated on the device as the reference does */
    float ortho_w, ortho_h;                   /* both > 0: the reference's ORTHO build (Renderer.cuh:13; getRayDirectionOrtho,
                                                 Renderer.cu:61-70,202-204) with SetOrthoWindowSize(ortho_w, ortho_h): every ray
                                                 runs along fwd, its origin moves across the window; 0: perspective */
} vx_camera;
typedef struct vx_env { float light_dir[3], light_color[3], ambient[3]; } vx_env; /* Renderer.cuh:33-37 */
typedef struct vx_render_opts {
    int32_t checkerboard;  /* 1 = reference (ENABLE_CHECKERBOARD_RENDER, Renderer.cu:5) */
    uint32_t frame_number; /* dFrameInfo.FrameNumber during the frame */
    int32_t debug_view;    /* 1 = shipped DEBUG_VIEW quadrants (Renderer.cu:4,215-243,270-275) */
    int32_t shadows;       /* 1 = sun shadow ray of Renderer.cu:102 enabled */
    int32_t crosshair;     /* 1 = reference (Renderer.cu:260-268) */
    int32_t norm_mode;     /* VX_NORM_* */
    uint32_t tile_x0, tile_y0, tile_w, tile_h; /* thread-space rectangle to render; w=h=0: whole frame */
    uint64_t* d_stats;     /* optional DEVICE pointer to 8 counters the frame adds to: primary rays, primary
                              hits, shadow rays, shadow hits, steps, C, K, E (work counters of vx_hits);
                              the counting variant of the kernel is slower — leave NULL when timing */
    uint32_t band_count, band_index; /* multi-GPU frame partition: the frame's 8-row bands are dealt round-robin
                              to band_count devices; this call renders the bands == band_index (mod band_count)
                              into their true positions of d_bgra.  band_count = 0: the whole frame */
} vx_render_opts;

void vx_get_directions(const float euler_xyz[3], float fwd[3], float up[3], float right[3]); /* Renderer.cu:27-42 */
/* d_bgra: W*H*4 bytes of device memory, BGRA8888 (Renderer.cuh:29-31); asynchronous on `stream`.
 * d_steps (optional, W*H int32): per-pixel step count incl. shadow steps.
 * ONE FRAME IN FLIGHT PER TRACER: a tracer owns the frame's hit queue, ray counters and timing events, so consecutive
 * vx_render / vx_bounce_rays / vx_trace_device calls on the same tracer must go to the same stream (or be separated by a
 * synchronisation); use one tracer per stream to overlap frames (tracers share the world, which is read-only). */
int vx_render(vx_tracer* t, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts,
              void* d_bgra, int32_t* d_steps, uint32_t W, uint32_t H, void* stream);
/* Same into host memory (render into the tracer's device frame, then D2H; synchronous): the
 * RenderScreen + cudaMemcpy pair of VoxelApp/main.cu:165-167. */
int vx_render_host(vx_tracer* t, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts,
                   void* h_bgra, uint32_t W, uint32_t H);

/* vx_render without any allocation.  NO COUNTERPART IN THE REFERENCE: added for callers
 * that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  It needs no
 * tracer: the frame's scratch (three 8-byte counters and the queue of primary hits waiting for their sun ray, 72 bytes per ray
 * id) is ONE file-scope device array of 256 + 72 * VX_STATIC_MAX_RAY_IDS bytes (18 MiB) that every GPU loading the library
 * holds until exit.  Same kernels, same pixels as vx_render(t, cam, env, opts, d_bgra, d_steps, W, H, stream).
 *   - limit: ray ids of the call = ceil(columns / 16) * (8-row bands of this call) * 128 <= VX_STATIC_MAX_RAY_IDS, i.e. frames up
 *     to 640x360 (tile_* / band_* fields of opts count only their share: larger frames can be rendered in pieces); beyond it the
 *     call returns VX_ERR_INVALID and vx_last_error() says how many ids were needed;
 *   - cam->tan_half must be > 0 (vx_device_tanf once at set-up; evaluating it per call would allocate) and opts->d_stats NULL;
 *   - the call itself runs cudaMemsetAsync (the counters) and up to four kernel launches on `stream`, nothing else: no cudaMalloc,
 *     cudaFree or allocating library routine.  The first call on a device looks the array's address up
 *     (cudaGetSymbolAddress) — make it once outside graph capture;
 *   - ALL calls on a device share the array, and it keeps the last call's contents (the call clears what must start at zero):
 *     calls that overlap on the GPU (two streams, two host threads, two captured graphs replayed at once) race; each process has its own copy.
 *     Order them on one stream or with events. */
#define VX_STATIC_MAX_RAY_IDS 262144ull
int vx_render_static(const vx_world* world, const vx_camera* cam, const vx_env* env, const vx_render_opts* opts,
                     void* d_bgra, int32_t* d_steps, uint32_t W, uint32_t H, void* stream);

/* Incoherent secondary rays (BASELINE config 5).  For every primary hit of the frame (cam, opts: checkerboard, frame_number,
 * tile/band fields and norm_mode are honoured) generates the ambient-occlusion sample ray of calculateColor's sample loop
 * (Renderer.cu:121-165, sample i = 0): seed = thread_y * W + thread_x + (frame_number + 1) * 1000, direction =
 * normalize(randomFloat(seed, seed*10, seed*100) * 2 - 1) reflected into the hemisphere of the shading normal, origin = hit
 * position + normal * 0.01.  Device outputs with room for W*H rays: origins xyz, directions xyz, pixel index y*W+x, and the
 * ray count (one u64).  Ray order follows the frame's hit queue and is not deterministic; `pixels` identifies the rays.
 * Trace them with vx_trace_device (the reference passes maxSteps = 8 here; config 5 uses the full cap). */
int vx_bounce_rays(vx_tracer* t, const vx_camera* cam, const vx_render_opts* opts, uint32_t W, uint32_t H,
                   float* d_origins_xyz, float* d_dirs_xyz, uint32_t* d_pixels, unsigned long long* d_count, void* stream);

/* Per-kernel device times of a frame.  A frame is up to four launches on `stream`: primary_rays (march of the
 * primary rays; misses and unshadowed modes write their pixels), then with shadows on shadow_setup, shadow_rays (march
 * of the sun rays of Renderer.cu:102) and shade_pixels.  After vx_tracer_set_timing(t, 1) every vx_render records CUDA
 * events around them; vx_tracer_frame_times waits for the last frame and returns the four durations in ms (0 for
 * launches the frame did not need). */
int vx_tracer_set_timing(vx_tracer* t, int enable);
/* Where the hit pixels are shaded: 0 (default) = in the dense shade_pixels pass after the sun-ray march, fastest when
 * the frame is in this GPU's memory; 1 = inside the march, 32 finished sun rays at a time (no shade_pixels launch):
 * the pixel stores are spread over the march, which is what a frame mapped from another GPU (vx_shared_open) wants —
 * partition.SharedFrame.attach sets it for such ranks.  Same pixels either way. */
int vx_tracer_set_fused_shade(vx_tracer* t, int mode);
int vx_tracer_frame_times(vx_tracer* t, float ms[4]);

/* Shared frame for the multi-GPU frame partition (no reference counterpart: main.cu:95-99 has one cudaMalloc'd
 * framebuffer on one GPU).  Rank 0 calls vx_shared_alloc and hands the 64-byte handle to the other ranks' processes,
 * which vx_shared_open it (CUDA IPC, peer access over NVLink / NVSwitch) and pass the mapped pointer as vx_render's
 * d_bgra together with their band_count / band_index: their kernels' pixel stores land in rank 0's HBM, and the only
 * exchange left is a completion barrier.  vx_device_to_host copies from any device pointer (e.g. that frame) and
 * waits for the copy. */
#define VX_SHARED_HANDLE_BYTES 64
int vx_shared_alloc(size_t bytes, int device, void** d_ptr, uint8_t handle[VX_SHARED_HANDLE_BYTES]);
int vx_shared_open(const uint8_t handle[VX_SHARED_HANDLE_BYTES], int device, void** d_ptr);
int vx_shared_close(void* d_ptr, int device); /* a pointer from vx_shared_open */
int vx_shared_free(void* d_ptr, int device);  /* a pointer from vx_shared_alloc */
int vx_device_to_host(void* h_dst, const void* d_src, size_t bytes, int device, void* stream);
int vx_device_fill(void* d_dst, int byte, size_t bytes, int device, void* stream); /* cudaMemset on any device pointer; waits */
/* The frame for a consumer on the host (main.cu:165-167 copies it down every frame) when several ranks render it: one
 * host frame of W*H*4 bytes in memory all the ranks' processes map (POSIX shared memory), page-locked in each process
 * by vx_host_register; after vx_render with band_count / band_index into its own device frame, each rank's
 * vx_bands_to_host enqueues the copy of exactly its 8-row bands into their places of the host frame on `stream` —
 * every rank down its own PCIe link, no GPU-to-GPU exchange.  The caller synchronises the streams and the ranks.
 * NOT for checkerboard frames: vx_render deals bands of 8 THREAD rows, and with checkerboard = 1 thread row ty writes
 * pixel rows 2 ty .. 2 ty + 2, so neighbouring bands interleave in pixel space and no row copy separates them.  Render
 * such partitions into one shared frame (vx_shared_alloc / vx_shared_open), where every pixel is stored by its owner. */
int vx_host_register(void* h_ptr, size_t bytes);
int vx_host_unregister(void* h_ptr);
int vx_bands_to_host(void* h_frame, const void* d_frame, uint32_t W, uint32_t H, uint32_t band_count, uint32_t band_index,
                     int device, void* stream);

/* Device math the parity harness shares with the CPU oracle. */
int vx_device_tanf(float x, float* out, int device);
int vx_device_normalize(const float* dirs_xyz, float* out_xyz, size_t n, int device); /* device rsqrtf */
/* Diagnostics: best-of-4 streaming read rate of a `bytes`-sized buffer read `reps` times by every SM
 * (ld.global.cg, 16 B per thread): the L2 roofline denominator when bytes << L2, HBM when bytes >> L2. */
int vx_measure_read_bandwidth(int device, size_t bytes, int reps, float* gb_per_s);

#ifdef __cplusplus
}
#endif
#endif
