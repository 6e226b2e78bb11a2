# Builds libgpuprobe.so (sm_100a only) in-tree and the CPU oracle used by the tests.
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := k8s-gpu-device-plugin_b200
CSRC := $(PKG)/csrc
LIB := $(PKG)/libgpuprobe.so
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
           -Xcompiler -fPIC,-fvisibility=hidden,-Wall -Xptxas -v
SRCS := $(CSRC)/engine.cu $(CSRC)/pattern.cu $(CSRC)/stream.cu $(CSRC)/gemm.cu $(CSRC)/p2p.cu $(CSRC)/noalloc.cu $(CSRC)/nvml_lite.cu $(CSRC)/arena.cu
OBJS := $(SRCS:.cu=.o)
HDRS := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/gpuprobe.h

HOST := $(PKG)/host
HOST_HDRS := $(wildcard $(HOST)/*.hpp) include/gpuprobe.h
CXX ?= g++
CXXFLAGS := -O2 -g -std=c++17 -Wall -Wextra -pthread -rdynamic

all: $(LIB) oracle host

host: $(HOST)/host_selftest $(HOST)/gpuprobe-slice $(HOST)/plugin_sim

$(HOST)/host_selftest: $(HOST)/host_selftest.cpp $(HOST_HDRS)
	$(CXX) $(CXXFLAGS) -o $@ $<

$(HOST)/gpuprobe-slice: $(HOST)/probe_slice.cpp $(LIB) include/gpuprobe.h
	$(CXX) $(CXXFLAGS) -o $@ $< -L$(PKG) -lgpuprobe -Wl,-rpath,'$$ORIGIN/..'

$(HOST)/plugin_sim: $(HOST)/plugin_sim.cpp $(LIB) $(HOST_HDRS)
	$(CXX) $(CXXFLAGS) -o $@ $< -L$(PKG) -lgpuprobe -Wl,-rpath,'$$ORIGIN/..'

$(CSRC)/%.o: $(CSRC)/%.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -c -o $@ $< 2> $@.ptxas.log || (cat $@.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS) -cudart static -lpthread -ldl

oracle:
	$(MAKE) -C oracle

# development build for tools/: the launch-shape / raster / cache-hint switches become environment variables
# (GPUPROBE_GEMM_GROUP_M, GPUPROBE_GEMM_HINTS, GPUPROBE_P2P_SHAPE).  The release library reads none of them.
DEVLIB := $(PKG)/libgpuprobe_dev.so
dev: $(DEVLIB)
$(DEVLIB): $(SRCS) $(HDRS)
	$(NVCC) $(NVFLAGS) -DGPUPROBE_DEV_SWITCHES -shared -o $@ $(SRCS) -cudart static -lpthread -ldl 2> $(CSRC)/dev.ptxas.log || (cat $(CSRC)/dev.ptxas.log; exit 1)

clean:
	rm -f $(OBJS) $(CSRC)/*.ptxas.log $(LIB) $(DEVLIB) $(HOST)/host_selftest $(HOST)/gpuprobe-slice $(HOST)/plugin_sim
	$(MAKE) -C oracle clean

# development sweeps (not part of `all`): launch shapes, peer-write store shapes, allocator timing
tools: tools/hbm_sweep tools/p2p_write_sweep tools/vmm_timing
tools/hbm_sweep: tools/hbm_sweep.cu
	$(NVCC) -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -lineinfo -o $@ $<
tools/p2p_write_sweep: tools/p2p_write_sweep.cu
	$(NVCC) -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o $@ $<
tools/vmm_timing: tools/vmm_timing.cu
	$(NVCC) -O2 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o $@ $< -L/usr/local/cuda/lib64/stubs -lcuda

.PHONY: all oracle host clean dev tools
