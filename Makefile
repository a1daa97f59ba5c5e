# Builds libvxrt.so (the product: CUDA kernels + C ABI) for sm_100a, in-tree.
NVCC   ?= nvcc
ARCH   := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := -std=c++17 -O3 $(ARCH) -lineinfo -fmad=false -Xcompiler -fPIC -Xptxas -v
CSRC   := voxelengine_b200/csrc
LIB    := voxelengine_b200/lib/libvxrt.so
SRCS   := $(CSRC)/vx_world.cu $(CSRC)/vx_trace.cu $(CSRC)/vx_shared.cu $(CSRC)/vx_edit.cu
HDRS   := $(CSRC)/vx_device.cuh $(CSRC)/vx_march.cuh $(CSRC)/vx_noise.cuh $(CSRC)/vx_host.h include/vxrt.h

all: $(LIB)

$(LIB): $(SRCS) $(HDRS)
	mkdir -p voxelengine_b200/lib
	$(NVCC) $(NVFLAGS) -shared -o $@ $(SRCS) 2> voxelengine_b200/lib/ptxas.log || (cat voxelengine_b200/lib/ptxas.log; exit 1)

oracle:
	$(MAKE) -C oracle all

clean:
	rm -f $(LIB) voxelengine_b200/lib/ptxas.log
.PHONY: all oracle clean
