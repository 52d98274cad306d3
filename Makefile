# Builds the C-ABI library of the B200 primary-ray path: librt_b200.so (sm_100a only, no CPU fallback).
PKG   := triangles_sdf_cpu_raytracing_b200
CSRC  := $(PKG)/csrc
NVCC  ?= /usr/local/cuda/bin/nvcc
ARCH  := -gencode arch=compute_100a,code=sm_100a
NVFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC -Xcompiler -Wall
OUT   := $(PKG)/librt_b200.so
OBJ   := $(CSRC)/build

HOST  := $(PKG)/hostapi
EXAMPLE := $(HOST)/example_frame_loop

PEAKS := tools/peaks

all: $(OUT) $(EXAMPLE) $(PEAKS)

# measured roofs (FP32 with / without FMA, issue rate, cached-gather rates, dependent-load latency): run once per box
$(PEAKS): tools/peaks.cu
	$(NVCC) $(ARCH) -O3 -std=c++17 -lineinfo -o $@ $<

$(OUT): $(OBJ)/rt_kernels_exact.o $(OBJ)/rt_kernels_fast.o $(OBJ)/rt_api.o $(OBJ)/bvh_builder.o $(OBJ)/bvh_gpu.o $(OBJ)/ingest.o $(OBJ)/rt_host.o
	$(NVCC) $(ARCH) -shared -o $@ $^ -cudart shared

# host mirror of the reference's C++ interface (pure host code; -ffp-contract=off keeps the camera matrices
# bit-identical to the reference's)
$(OBJ)/rt_host.o: $(HOST)/rt_host.cpp $(HOST)/scenes.hpp $(HOST)/raytracing.hpp $(HOST)/camera.hpp $(HOST)/rt_math.hpp include/rt_b200.h include/rt_b200_host.h
	@mkdir -p $(OBJ)
	g++ -std=c++17 -O2 -ffp-contract=off -fPIC -Wall -Wextra -c $< -o $@

$(EXAMPLE): $(HOST)/example_frame_loop.cpp $(HOST)/scenes.hpp $(HOST)/raytracing.hpp $(HOST)/camera.hpp $(HOST)/rt_math.hpp $(OUT)
	g++ -std=c++17 -O2 -ffp-contract=off -Wall -Wextra -o $@ $< -L$(PKG) -lrt_b200 -Wl,-rpath,'$$ORIGIN/..'

$(OBJ)/rt_kernels_exact.o: $(CSRC)/rt_kernels_exact.cu $(CSRC)/rt_kernels.cuh $(CSRC)/rt_shade.cuh $(CSRC)/rt_device.cuh $(CSRC)/rt_scene.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -fmad=false -c $< -o $@

$(OBJ)/rt_kernels_fast.o: $(CSRC)/rt_kernels_fast.cu $(CSRC)/rt_kernels.cuh $(CSRC)/rt_shade.cuh $(CSRC)/rt_device.cuh $(CSRC)/rt_scene.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -fmad=true -c $< -o $@

$(OBJ)/rt_api.o: $(CSRC)/rt_api.cu $(CSRC)/rt_scene.h $(CSRC)/bvh_builder.h $(CSRC)/ingest.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(OBJ)/bvh_gpu.o: $(CSRC)/bvh_gpu.cu $(CSRC)/bvh_builder.h $(CSRC)/rt_scene.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(OBJ)/ingest.o: $(CSRC)/ingest.cu $(CSRC)/ingest.h $(CSRC)/rt_scene.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(OBJ)/bvh_builder.o: $(CSRC)/bvh_builder.cpp $(CSRC)/bvh_builder.h $(CSRC)/rt_scene.h include/rt_b200.h
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

.PHONY: clean
clean:
	rm -rf $(OBJ) $(OUT) $(EXAMPLE) $(PEAKS)
