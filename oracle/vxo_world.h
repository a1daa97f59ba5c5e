/* vxo_world.h — CPU ORACLE world container (test infrastructure, not product code). */
#ifndef VXO_WORLD_H
#define VXO_WORLD_H
#include <cstdint>
#include <vector>

struct vxo_world {
    uint32_t X = 0, Y = 0, Z = 0;    /* voxels */
    uint32_t cx = 0, cy = 0, cz = 0; /* bricks  */
    int factor = 8;
    size_t words_per_brick = 16;
    std::vector<uint32_t> dense;      /* reference dense grid, tiled index (R1/R2) */
    std::vector<uint32_t> top;        /* coarse occupancy bits, tiled index over (cx,cy,cz) */
    std::vector<uint32_t> brick_bits; /* per brick factor^3 bits, in-brick tiled index */
    std::vector<float> bounds;        /* per brick min xyz, max xyz (inclusive voxel coords) */
    /* what the march reads: the vectors above, or caller-owned arrays (vxo_world_from_brickmap_view) */
    const uint32_t* top_p = nullptr;
    const uint32_t* brick_p = nullptr;
    const float* bounds_p = nullptr;
    void bind() { top_p = top.data(); brick_p = brick_bits.data(); bounds_p = bounds.data(); }
};
#endif
