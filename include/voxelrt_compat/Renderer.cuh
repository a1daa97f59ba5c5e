// Renderer.cuh — drop-in replacement for the reference header VoxelRT/Renderer.cuh (see VoxelRT.h).
#pragma once
#include "VoxelRT.h"
