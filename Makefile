# wave3d-b200 build. Knobs mirror the reference Makefile (ref: Makefile:1-4):
#   ARCH       GPU target (default sm_100a; the code uses TMA/mbarrier and needs sm_90+)
#   HOST_COMP  host C++ compiler
ARCH      ?= sm_100a
HOST_COMP ?= g++
NVCC      ?= /usr/local/cuda/bin/nvcc
CUDA_INC  ?= /usr/local/cuda/include

COMPUTE   := $(subst sm_,compute_,$(ARCH))
NVFLAGS   := -gencode arch=$(COMPUTE),code=$(ARCH) -O3 -std=c++17 -lineinfo -Xcompiler -fPIC,-fopenmp -Iinclude -Icsrc
CXXFLAGS  := -O3 -std=c++17 -fPIC -fopenmp -ffp-contract=off -Wall -Wextra -Iinclude -Icsrc

OBJ := build/wave3d_spec.o build/wave3d_cpu.o build/wave3d_kernels.o build/wave3d_runtime.o

all: lib/libwave3d.so bin/wave3d

build/%.o: csrc/%.cpp include/wave3d.h
	@mkdir -p build
	$(HOST_COMP) $(CXXFLAGS) -c $< -o $@

build/%.o: csrc/%.cu include/wave3d.h csrc/wave3d_internal.hpp $(wildcard csrc/*.inc)
	@mkdir -p build
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@ 2> build/$*.ptxas.log || (cat build/$*.ptxas.log; false)

lib/libwave3d.so: $(OBJ)
	@mkdir -p lib
	$(NVCC) -shared -o $@ $(OBJ) -Xcompiler -fopenmp -lgomp

bin/wave3d: csrc/wave3d_cli.cpp lib/libwave3d.so
	@mkdir -p bin
	$(HOST_COMP) -O2 -Iinclude $< -o $@ -Llib -lwave3d -Wl,-rpath,'$$ORIGIN/../lib'

sass: lib/libwave3d.so
	@mkdir -p profiles
	/usr/local/cuda/bin/cuobjdump -sass build/wave3d_kernels.o > profiles/wave3d_kernels.sass

clean:
	rm -rf build lib bin

.PHONY: all clean sass
