#!/bin/bash
# Builds tuning variants of libvxrt.so: tools/build_variants.sh name "-DFLAG=.. -DFLAG=.." [name flags]...
# Output: voxelengine_b200/lib/variants/libvxrt_<name>.so (git-ignored, travels with gpurun); run with VXRT_LIB=<path>.
set -e
cd "$(dirname "$0")/.."
mkdir -p voxelengine_b200/lib/variants
while [ $# -gt 1 ]; do
  name=$1; flags=$2; shift 2
  nvcc -std=c++17 -O3 -gencode arch=compute_100a,code=sm_100a -lineinfo -fmad=false -Xcompiler -fPIC $flags -shared \
    -o voxelengine_b200/lib/variants/libvxrt_$name.so voxelengine_b200/csrc/vx_world.cu voxelengine_b200/csrc/vx_trace.cu voxelengine_b200/csrc/vx_shared.cu voxelengine_b200/csrc/vx_edit.cu &
done
wait
ls -la voxelengine_b200/lib/variants/
