# Builds libgpuprobe.so (sm_100a only) in-tree and the CPU oracle used by the tests.
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := k8s-gpu-device-plugin_b200
CSRC := $(PKG)/csrc
LIB := $(PKG)/libgpuprobe.so
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
           -Xcompiler -fPIC,-fvisibility=hidden,-Wall -Xptxas -v
SRCS := $(CSRC)/engine.cu $(CSRC)/pattern.cu $(CSRC)/stream.cu $(CSRC)/gemm.cu $(CSRC)/p2p.cu $(CSRC)/noalloc.cu
OBJS := $(SRCS:.cu=.o)
HDRS := $(wildcard $(CSRC)/*.h $(CSRC)/*.cuh) include/gpuprobe.h

all: $(LIB) oracle

$(CSRC)/%.o: $(CSRC)/%.cu $(HDRS)
	$(NVCC) $(NVFLAGS) -c -o $@ $< 2> $@.ptxas.log || (cat $@.ptxas.log; exit 1)

$(LIB): $(OBJS)
	$(NVCC) -shared -o $@ $(OBJS) -cudart static -lpthread -ldl

oracle:
	$(MAKE) -C oracle

clean:
	rm -f $(OBJS) $(CSRC)/*.ptxas.log $(LIB)
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
