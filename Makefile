# Builds libperch_b200.so (product, sm_100a only) and the test oracles.
NVCC      ?= nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
CSRC      := perception_b200/csrc
NVFLAGS   := -std=c++17 -O3 $(ARCH) -lineinfo -Xcompiler -fPIC,-ffp-contract=off -I$(CSRC) -Iinclude
SRCS      := $(CSRC)/pose_kernels.cu $(CSRC)/cloud.cu $(CSRC)/colour.cu $(CSRC)/knn.cu $(CSRC)/gicp.cu $(CSRC)/icp.cu $(CSRC)/states.cu $(CSRC)/input.cu $(CSRC)/compose.cu $(CSRC)/raster_full.cu $(CSRC)/verify.cu $(CSRC)/obs_grid.cu $(CSRC)/topk.cu $(CSRC)/capi.cu $(CSRC)/capi_multi.cu $(CSRC)/capi_states.cu $(CSRC)/capi_input.cu $(CSRC)/capi_compose.cu
OBJS      := $(SRCS:.cu=.o) $(CSRC)/clusters.o
LIB       := perception_b200/libperch_b200.so
REF       := /root/reference

all: $(LIB) oracle

$(CSRC)/%.o: $(CSRC)/%.cu $(wildcard $(CSRC)/*.cuh $(CSRC)/*.hpp) include/perch_b200.h
	$(NVCC) $(NVFLAGS) -c $< -o $@

$(CSRC)/clusters.o: $(CSRC)/clusters.cpp $(CSRC)/clusters.hpp
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS) -lnccl

# Debug variant with per-level survivor counters in the per-pose kernel (tools/pose_stats.py); never the product.
# A/B build of the per-pose kernel with extra -D switches: make variant VARIANT_FLAGS=-DFOO VARIANT=foo
#   -> perception_b200/_ab/libperch_$(VARIANT).so (tools/ab_cost.py)
variant: $(OBJS)
	mkdir -p perception_b200/_ab
	$(NVCC) $(NVFLAGS) $(VARIANT_FLAGS) -c $(CSRC)/pose_kernels.cu -o $(CSRC)/pose_kernels_variant.o
	$(NVCC) $(ARCH) -shared -o perception_b200/_ab/libperch_$(VARIANT).so $(CSRC)/pose_kernels_variant.o $(filter-out $(CSRC)/pose_kernels.o,$(OBJS)) -lnccl

stats: $(OBJS)
	$(NVCC) $(NVFLAGS) -DPERCH_STATS -c $(CSRC)/pose_kernels.cu -o $(CSRC)/pose_kernels_stats.o
	$(NVCC) $(ARCH) -shared -o perception_b200/libperch_b200_stats.so $(CSRC)/pose_kernels_stats.o $(filter-out $(CSRC)/pose_kernels.o,$(OBJS)) -lnccl

# Bounds-checked build of the per-pose kernel (tools/bounds_check.py): compute-sanitizer is closed on this pool.
check: $(OBJS)
	$(NVCC) $(NVFLAGS) -DPERCH_BOUNDS_CHECK -c $(CSRC)/pose_kernels.cu -o $(CSRC)/pose_kernels_check.o
	$(NVCC) $(ARCH) -shared -o perception_b200/libperch_b200_check.so $(CSRC)/pose_kernels_check.o $(filter-out $(CSRC)/pose_kernels.o,$(OBJS)) -lnccl

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(OBJS) $(LIB)
	$(MAKE) -C oracle clean

.PHONY: all oracle clean stats check variant
