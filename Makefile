# libfedb200.so: every CUDA/C++ source of the framework, sm_100a only, in-tree (the built .so
# travels to the GPU box with the repo snapshot; it is git-ignored).
NVCC      ?= /usr/local/cuda/bin/nvcc
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := -O3 -std=c++17 -lineinfo $(ARCH) -Xcompiler -fPIC -Iinclude --expt-relaxed-constexpr \
             -Xptxas -v -cudart static
SRCS      := $(wildcard csrc/*.cu)
OBJS      := $(patsubst csrc/%.cu,build/%.o,$(SRCS))
LIB       := fedb200/lib/libfedb200.so

all: $(LIB)

build/%.o: csrc/%.cu $(wildcard csrc/*.cuh) include/fedb200.h
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; exit 1)

$(LIB): $(OBJS)
	@mkdir -p fedb200/lib
	$(NVCC) -shared $(ARCH) -cudart static -o $@ $(OBJS) -lpthread -ldl -lrt

clean:
	rm -rf build $(LIB)

.PHONY: all clean
