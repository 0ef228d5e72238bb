# Native build of the b200dp runtime + kernels (C ABI, loaded through ctypes).
# Cross-compiles for sm_100a without a GPU; the .so files stay in-tree so the
# gpurun snapshot carries them to the B200 box.
NVCC      ?= /usr/local/cuda/bin/nvcc
CXX       ?= g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden --expt-relaxed-constexpr
CXXFLAGS  := -O2 -std=c++17 -fPIC -fvisibility=hidden -Wall -pthread
PKG       := torch_distributed_sandbox_b200
CUDA_SRCS := $(wildcard csrc/*.cu)
CUDA_OBJS := $(patsubst csrc/%.cu,build/%.o,$(CUDA_SRCS))
HOST_SRCS := $(wildcard csrc/*.cpp)
HOST_OBJS := $(patsubst csrc/%.cpp,build/%.host.o,$(HOST_SRCS))

all: $(PKG)/lib/libb200dp_cuda.so $(PKG)/lib/libb200dp_host.so

build/%.o: csrc/%.cu csrc/common.h csrc/ptx.cuh
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

build/%.host.o: csrc/%.cpp
	@mkdir -p build
	$(CXX) $(CXXFLAGS) -c $< -o $@

$(PKG)/lib/libb200dp_cuda.so: $(CUDA_OBJS)
	@mkdir -p $(PKG)/lib
	$(NVCC) $(ARCH) -shared -o $@ $^

$(PKG)/lib/libb200dp_host.so: $(HOST_OBJS)
	@mkdir -p $(PKG)/lib
	$(CXX) -shared -pthread -o $@ $^ -lrt

clean:
	rm -rf build $(PKG)/lib/*.so

.PHONY: all clean
