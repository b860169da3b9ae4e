# Builds libb200fem.so (hand-written sm_100a kernels + C ABI) and the CPU oracle (test infrastructure).
NVCC      ?= /usr/local/cuda/bin/nvcc
HOSTCXX   ?= /usr/bin/g++
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVCCFLAGS := $(ARCH) -O3 -std=c++17 -lineinfo -ccbin $(HOSTCXX) -Xcompiler -fPIC,-Wall -Iinclude -Iarcanefem_b200/csrc $(EXTRA_NVCCFLAGS)
SRC       := $(wildcard arcanefem_b200/csrc/*.cu)
OBJ       := $(patsubst arcanefem_b200/csrc/%.cu,build/%.o,$(SRC))
LIB       := arcanefem_b200/lib/libb200fem.so
HDR       := $(wildcard include/*.h include/b2f/*.h include/b2f/*.cuh arcanefem_b200/csrc/*.h arcanefem_b200/csrc/*.cuh)

all: $(LIB) oracle

$(LIB): $(OBJ)
	@mkdir -p $(dir $@)
	$(NVCC) $(ARCH) -shared -ccbin $(HOSTCXX) -o $@ $(OBJ) -ldl

build/%.o: arcanefem_b200/csrc/%.cu $(HDR)
	@mkdir -p build
	$(NVCC) $(NVCCFLAGS) -c $< -o $@

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf build $(LIB) oracle/liboracle.so

.PHONY: all oracle clean
