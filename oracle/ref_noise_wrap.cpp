/* C entry points around the REFERENCE's cuda_noise.cuh compiled as plain C++ (test infrastructure,
 * built only into oracle/_ref/).  float->unsigned conversions follow x86-64 here, not CUDA. */
#include "cuda_noise.cuh"
namespace cudaNoise { float gradMap[16][3] = {}; }
extern "C" unsigned vxref_hash(unsigned s) { return cudaNoise::hash(s); }
extern "C" float vxref_random_float(unsigned s) { return cudaNoise::randomFloat(s); }
extern "C" float vxref_perlin(float x, float y, float z, float scale, int seed) { return cudaNoise::perlinNoise(make_float3(x, y, z), scale, seed); }
extern "C" float vxref_terrain_noise(float x, float y, float z) {
    /* VoxelWorldBuilder.cu:4-8 */
    return cudaNoise::repeaterPerlin(make_float3(x, y, z), 1.0f, 0x71889283, 32, 2.0f, 0.5f);
}
extern "C" float vxref_fbm(float x, float y, float z, int octaves) {
    return cudaNoise::repeaterPerlin(make_float3(x, y, z), 1.0f, 0x71889283, octaves, 2.0f, 0.5f);
}
