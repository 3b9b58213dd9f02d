# libscenesubmit — build everything in-tree (the .so files travel to the GPU box with the snapshot).
#   make            product library  vulkanengine_b200/libscenesubmit.so   (nvcc, sm_100a only)
#   make oracle     CPU oracle + reference-arithmetic build (test infrastructure)
#   make host       C++ host mirror demo/test binary
NVCC     ?= /usr/local/cuda/bin/nvcc
ARCH     := -gencode arch=compute_100a,code=sm_100a
# -fmad=false: no FMA contraction anywhere, the float op order is the reference's
NVFLAGS  := $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false $(EXTRA) \
            -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math,-fvisibility=hidden,-Wall
CSRC     := vulkanengine_b200/csrc
OBJDIR   := build
LIB      := vulkanengine_b200/libscenesubmit.so
CU       := $(CSRC)/transform_cull.cu $(CSRC)/transform_cull_x2.cu $(CSRC)/radix_sort.cu $(CSRC)/emit.cu $(CSRC)/resident.cu $(CSRC)/merge.cu $(CSRC)/raster.cu $(CSRC)/context.cu
CPP      := $(CSRC)/host_views.cpp
OBJS     := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU)) $(patsubst $(CSRC)/%.cpp,$(OBJDIR)/%.o,$(CPP))
HDRS     := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh include/*.h)

HOSTBIN  := vulkanengine_b200/host/host_cli.bin
HOSTSRC  := vulkanengine_b200/host/Scene.cpp vulkanengine_b200/host/SceneSerializer.cpp vulkanengine_b200/host/MeshImporter.cpp vulkanengine_b200/host/host_cli.cpp
HOSTHDR  := $(wildcard vulkanengine_b200/host/*.h) include/scenesubmit.h

MERGEBIN := vulkanengine_b200/host/merge_cli.bin

all: $(LIB) $(HOSTBIN) $(MERGEBIN)

# C++ host mirror of Scene / SceneRenderer: plain g++, links the C ABI only
$(HOSTBIN): $(HOSTSRC) $(HOSTHDR) $(LIB)
	g++ -std=c++17 -O2 -Wall -ffp-contract=off -o $@ $(HOSTSRC) -Lvulkanengine_b200 -lscenesubmit -lz \
	    -Wl,-rpath,'$$ORIGIN/..'

# merged draw lists of several shards from a C++ host (C ABI only)
$(MERGEBIN): vulkanengine_b200/host/merge_cli.cpp include/scenesubmit.h $(LIB)
	g++ -std=c++17 -O2 -Wall -o $@ vulkanengine_b200/host/merge_cli.cpp -Lvulkanengine_b200 -lscenesubmit -Wl,-rpath,'$$ORIGIN/..'

host: $(HOSTBIN) $(MERGEBIN)

$(OBJDIR)/%.o: $(CSRC)/%.cu $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; false)

$(OBJDIR)/%.o: $(CSRC)/%.cpp $(HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -o $@ $(OBJS)

oracle:
	$(MAKE) -C oracle all

clean:
	rm -rf $(OBJDIR) $(LIB)
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
