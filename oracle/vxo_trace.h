/* vxo_trace.h — CPU ORACLE internals shared by vxo_trace.cpp and vxo_render.cpp (test infrastructure). */
#ifndef VXO_TRACE_H
#define VXO_TRACE_H
#include "vxo_common.h"
struct vxo_world;

namespace vxo {

struct Level { const uint32_t* bits; int nx, ny, nz; };
struct Work { uint32_t C = 0, K = 0, E = 0; };
struct DdaOut {
    bool hit, oob;
    int cell[3];  /* HitCell */
    int next[3];  /* NextCell */
    V3 point;     /* HitIntersectedPoint */
    V3 normal;    /* HitNormal */
    int steps;    /* stepsTaken */
};
struct MarchOut {
    bool hit = false;
    int steps = 0, coarse_steps = 0, fine_steps = 0;
    V3 pos{0, 0, 0}, normal{0, 0, 0};
    int voxel[3] = {-1, -1, -1};
    uint32_t brick = 0xFFFFFFFFu;
    Work work;
};

bool ray_aabb(V3 s, V3 d, V3 bmin, V3 bmax, V3* point, V3* normal);
void dda(const Level& L, V3 s, V3 d, const float* bounds6, int scale, int clip_max, DdaOut& r, Work* wk);
bool march(const vxo_world& w, V3 origin, V3 ray, int norm_mode, int max_steps, MarchOut& o);

} // namespace vxo
#endif
