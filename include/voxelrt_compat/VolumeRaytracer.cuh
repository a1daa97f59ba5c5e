// VolumeRaytracer.cuh — drop-in replacement for the reference header VoxelRT/VolumeRaytracer.cuh (see VoxelRT.h).
#pragma once
#include "VoxelRT.h"
