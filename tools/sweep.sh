#!/bin/bash
# Runs tools/prof_frame.py for every tuning variant (see build_variants.sh): tools/sweep.sh [workload] [frames]
cd "$(dirname "$0")/.."
for so in voxelengine_b200/lib/variants/libvxrt_*.so; do
  echo "== $so"; VXRT_LIB=$PWD/$so timeout 300 python tools/prof_frame.py ${1:-w2-4k} ${2:-5} 2>&1 | tail -1
done
