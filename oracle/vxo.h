/*
 * vxo.h — C interface of the CPU ORACLE (test infrastructure, NOT product code).
 *
 * The oracle is a CPU restatement of the reference's hot path
 * (JoshuaLim007/VoxelEngine, VoxelRT/): the bit-array voxel storage and tiled
 * index, the noise-terrain world build, the host brickmap build, the two-level
 * DDA ray march, and the screen renderer's pixel logic.  Only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs
 * may load it.  The product (voxelengine_b200/) never links or calls it.
 *
 * Numerical definition restated: IEEE-754 binary32, round to
 * nearest, no FMA contraction, CUDA's saturating float->int / float->uint
 * conversions.  normalize() is selectable: VXO_NORM_HOST restates the
 * reference's host definition of rsqrtf (CUDA crt/math_functions.hpp host
 * path: (float)(1.0/sqrt((double)x))); VXO_NORM_NONE takes directions that the
 * caller already normalised (e.g. on the GPU with the device rsqrtf).
 */
#ifndef VXO_H
#define VXO_H
#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

enum { VXO_NORM_HOST = 1, VXO_NORM_NONE = 2 };
/* float->unsigned conversion used by the noise hash (cuda_noise.cuh:118-121):
 * CUDA saturates (cvt.rzi.u32.f32); x86-64 g++ wraps modulo 2^32 (cvttss2si r64). */
enum { VXO_F2U_CUDA = 0, VXO_F2U_X86 = 1 };

typedef struct vxo_world vxo_world;

/* ---- index helpers (VolumeRaytracer.cuh:107-171, widened to 64 bit) ---- */
uint64_t vxo_sample_index(uint32_t x, uint32_t y, uint32_t z, uint32_t width, uint32_t height);
void vxo_position_from_index(uint64_t index, uint32_t width, uint32_t height,
                             uint32_t* x, uint32_t* y, uint32_t* z);

/* ---- noise (cuda_noise.cuh:44-54,118-121,162-202,565-628; VoxelWorldBuilder.cu:4-35) ---- */
uint32_t vxo_hash(uint32_t seed);
float vxo_perlin(float px, float py, float pz, float scale, int seed, int f2u_mode);
float vxo_terrain_noise(float fx, float fy, float fz, int f2u_mode); /* PerlinNoise() */
int vxo_voxel_solid(uint32_t x, uint32_t y, uint32_t z, int f2u_mode); /* PopulateVoxels body */
/* dense tiled bit grid of a window [x0,x0+X) x [y0,..) x [z0,..) of the infinite noise field;
 * the window is indexed locally (GetSampleIndex(x-x0, y-y0, z-z0, X, Y)). */
void vxo_populate_dense(uint32_t* bits, uint32_t X, uint32_t Y, uint32_t Z,
                        uint32_t x0, uint32_t y0, uint32_t z0, int f2u_mode, int nthreads);

/* ---- brickmap (VolumeRaytracer.cuh:379-516) ---- */
vxo_world* vxo_world_from_dense(const uint32_t* tiled_bits, uint32_t X, uint32_t Y, uint32_t Z, int factor);
vxo_world* vxo_world_from_noise(uint32_t X, uint32_t Y, uint32_t Z, int factor, int f2u_mode, int nthreads);
/* non-owning view of an existing reference-form brickmap (no dense grid): for the timed CPU baseline */
vxo_world* vxo_world_from_brickmap_view(const uint32_t* top_bits, const uint32_t* brick_bits, const float* bounds6,
                                        uint32_t X, uint32_t Y, uint32_t Z, int factor);
void vxo_world_free(vxo_world* w);
void vxo_world_dims(const vxo_world* w, uint32_t dims_xyz[3], uint32_t coarse_xyz[3], int* factor);
const uint32_t* vxo_world_dense_bits(const vxo_world* w);   /* X*Y*Z bits, tiled index */
const uint32_t* vxo_world_top_bits(const vxo_world* w);     /* coarse bits, tiled index  */
const uint32_t* vxo_world_brick_bits(const vxo_world* w);   /* nbricks * factor^3/32 words, brick-major by tiled coarse index */
const float* vxo_world_bounds(const vxo_world* w);          /* nbricks * 6 floats (min xyz, max xyz) */

/* ---- slab test and one-level DDA, exposed for unit tests (VolumeRaytracer.cu:124-352) ---- */
int vxo_ray_aabb(const float start[3], const float dir[3], const float bmin[3], const float bmax[3],
                 float out_point[3], float out_normal[3]);

/* ---- two-level march (VolumeRaytracer.cu:354-525) + batch post-processing (:601-614) ---- */
typedef struct vxo_hits {
    uint8_t* hit;        /* n      */
    float* pos;          /* n*3, +inf on miss (dispatch kernel, VolumeRaytracer.cu:108-113) */
    float* normal;       /* n*3, un-negated (points along the ray) */
    int32_t* steps;      /* n      */
    int32_t* voxel;      /* n*3  hit voxel = coarse cell*factor + fine cell; -1 on miss */
    uint32_t* brick;     /* n    tiled coarse index of the brick that was hit; 0xFFFFFFFF on miss */
    int32_t* coarse_steps; /* n */
    int32_t* fine_steps;   /* n */
    uint32_t* counters;  /* n*3: C coarse cells examined, K occupied cells examined, E bricks entered */
} vxo_hits;              /* any pointer may be NULL */

void vxo_trace(const vxo_world* w, const float* origins, const float* dirs, size_t n,
               int norm_mode, int max_steps, const vxo_hits* out, int nthreads);

/* single-level DDA on the dense grid (config 1 (i)): returns hit + steps only */
void vxo_trace_dense(const vxo_world* w, const float* origins, const float* dirs, size_t n,
                     int norm_mode, uint8_t* hit, int32_t* steps, int nthreads);

/* ---- screen renderer (Renderer.cu:27-276) ---- */
typedef struct vxo_camera {
    float origin[3], fwd[3], up[3], right[3];
    float fov_deg;
    float tan_half; /* > 0: use this value for tanf(fov/2) (e.g. the device's), else host tanf */
    float ortho_w, ortho_h; /* both > 0: the ORTHO build (Renderer.cuh:13, Renderer.cu:61-70,202-204) with this OrthoSize */
} vxo_camera;
typedef struct vxo_env { float light_dir[3], light_color[3], ambient[3]; } vxo_env;
typedef struct vxo_render_opts {
    int checkerboard;   /* 1 = reference behaviour (Renderer.cu:5,186-194) */
    uint32_t frame_number; /* value dFrameInfo.FrameNumber holds during the frame */
    int debug_view;     /* 1 = shipped DEBUG_VIEW quadrants, 0 = calculateColor+Tonemap */
    int shadows;        /* 1 = shadow Raytrace restored (Renderer.cu:102), 0 = as shipped (hit=false) */
    int crosshair;      /* 1 = reference (Renderer.cu:260-268) */
    int norm_mode;
} vxo_render_opts;

void vxo_get_directions(const float euler[3], float fwd[3], float up[3], float right[3]);
/* per-pixel primary direction BEFORE Raytrace's own normalize; tan_half = tanf(fov/2) passed in so
 * CPU and GPU can share one value (H1 iii). */
void vxo_render(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                uint32_t W, uint32_t H, uint8_t* bgra, int32_t* steps_plane, int nthreads);

/* bounded sample of a frame (every ty_stride-th thread row in [ty_begin, ty_end)) for the timed CPU
 * baseline; bgra may be NULL; stats_out[8] = primary rays, primary hits, shadow rays, shadow hits,
 * steps, C, K, E */
void vxo_render_sample(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                       uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                       uint8_t* bgra, uint64_t* stats_out, int nthreads);

/* the same sample with its pixels (W*H*4 bytes, unsampled rows untouched) and step plane (W*H int32) written */
void vxo_render_sample_planes(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                              uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                              uint8_t* bgra, int32_t* steps_plane, int nthreads);
/* the rays screenDispatch gives Raytrace on those rows (primary rays, and a sun ray per primary hit), as a batch */
size_t vxo_frame_rays(const vxo_world* w, const vxo_camera* cam, const vxo_env* env, const vxo_render_opts* o,
                      uint32_t W, uint32_t H, uint32_t ty_begin, uint32_t ty_end, uint32_t ty_stride,
                      float* origins, float* dirs, size_t capacity);

#ifdef __cplusplus
}
#endif
#endif
