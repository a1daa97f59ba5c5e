// VoxelWorldBuilder.cuh — drop-in replacement for the reference header VoxelRT/VoxelWorldBuilder.cuh (see VoxelRT.h).
#pragma once
#include "VoxelRT.h"
using namespace GPUDDA;
