# Build of the native components (sm_100a only).
# Counterpart of the reference's 29-line Makefile (readme.md:52-80), which built the
# single file mpigpu-1.cu with `nvcc -O3 -std=c++11 -arch=sm_60` against Spectrum MPI.
NVCC      ?= nvcc
# the image exports CXX=/opt/gcc/bin/g++, a wrapper without libgomp; use the system compiler
CXX       := $(firstword $(wildcard /usr/bin/g++) g++)
ARCH      ?= -gencode arch=compute_100a,code=sm_100a
OUT       := mpi_cuda_b200/_native
BIN       := bin
CXXFLAGS  := -O3 -std=c++17 -fPIC -fopenmp -march=x86-64-v3 -ffp-contract=off -Iinclude -Wall -Wno-unknown-pragmas
NVCCFLAGS := -ccbin $(CXX) $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false -Iinclude -Xcompiler -fPIC,-fopenmp
CU_SRCS   := $(wildcard csrc/cuda/*.cu)
CU_HDRS   := $(wildcard csrc/cuda/*.cuh) $(wildcard include/*.h)

all: cpu cuda tools

cpu: $(OUT)/libwave3d_cpu.so $(BIN)/wave $(BIN)/openmpwave $(BIN)/mpi $(BIN)/mpiomp

$(OUT)/libwave3d_cpu.so: csrc/cpu/wave_cpu.cpp include/wave3d_cpu.h include/stencil_core.h
	@mkdir -p $(OUT)
	$(CXX) $(CXXFLAGS) -shared -o $@ csrc/cpu/wave_cpu.cpp -lpthread

$(BIN)/wave: csrc/cpu/wave_cli.cpp csrc/cpu/wave_cpu.cpp include/wave3d_cpu.h include/stencil_core.h
	@mkdir -p $(BIN)
	$(CXX) $(CXXFLAGS) -o $@ csrc/cpu/wave_cli.cpp csrc/cpu/wave_cpu.cpp -lpthread

$(BIN)/openmpwave $(BIN)/mpi $(BIN)/mpiomp: $(BIN)/wave
	cp $(BIN)/wave $@

cuda: $(OUT)/libwave3d_b200.so

$(OUT)/libwave3d_b200.so: $(CU_SRCS) $(CU_HDRS)
	@mkdir -p $(OUT)
	$(NVCC) $(NVCCFLAGS) -shared -o $@ $(CU_SRCS) -cudart static

tools: $(BIN)/mpigpu

# not part of `make all`: `make probe`
probe: $(BIN)/tma_probe3

# bring-up probe that established the 16-byte TMA coordinate rule
$(BIN)/tma_probe3: csrc/tools/tma_probe3.cu csrc/cuda/ptx.cuh
	@mkdir -p $(BIN)
	$(NVCC) -ccbin $(CXX) $(ARCH) -O2 -std=c++17 -Iinclude -o $@ csrc/tools/tma_probe3.cu

$(BIN)/mpigpu: csrc/tools/mpigpu_main.cpp $(OUT)/libwave3d_b200.so
	@mkdir -p $(BIN)
	$(NVCC) -ccbin $(CXX) $(ARCH) -O3 -std=c++17 -Iinclude -o $@ csrc/tools/mpigpu_main.cpp -L$(OUT) -lwave3d_b200 -Xlinker -rpath,'$$ORIGIN/../$(OUT)' -Xcompiler -pthread

sass: $(OUT)/libwave3d_b200.so
	@mkdir -p profiles/sass
	cuobjdump -sass $(OUT)/libwave3d_b200.so > profiles/sass/libwave3d_b200.sass

clean:
	rm -rf $(OUT)/*.so $(BIN) build

.PHONY: all cpu cuda tools probe sass clean
