# Builds the product libraries in-tree (they travel to the GPU box with the gpurun snapshot):
#   omp-bowtie2-prime_b200/libbt2gpu.so   CUDA kernels + C-ABI (sm_100a only)
#   omp-bowtie2-prime_b200/libbt2host.so  host-side index loader / FASTQ reader / SAM sink used by tests, bench and the CLI
# and the test-only oracle (oracle/liboracle.so, oracle/oracle_align, oracle/_ref/*).
PKG   := omp-bowtie2-prime_b200
GPU   := $(PKG)/csrc/gpu
HOST  := $(PKG)/csrc/host
NVCC  ?= nvcc
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -Xcompiler -fPIC -shared -Xcompiler -pthread -Iinclude
GCCLIB := $(dir $(shell g++ -print-libgcc-file-name))
CXXFLAGS := -O2 -std=c++17 -fPIC -fopenmp -B$(GCCLIB) -L$(GCCLIB) -Iinclude -w

# the long-read build: the same sources with the reference's two size constants raised, as a second library
LONGDEF := -DBT2GPU_MAX_ROWS=1000 -DBT2GPU_MAX_COLS=1072

all: $(PKG)/libbt2gpu.so $(PKG)/libbt2gpu_long.so $(PKG)/libbt2host.so oracle

$(PKG)/libbt2gpu_long.so: $(wildcard $(GPU)/*.cu $(GPU)/*.cuh) include/bt2gpu.h
	$(NVCC) $(NVFLAGS) $(LONGDEF) -o $@ $(GPU)/bt2gpu.cu

$(PKG)/libbt2gpu.so: $(wildcard $(GPU)/*.cu $(GPU)/*.cuh) include/bt2gpu.h
	$(NVCC) $(NVFLAGS) -o $@ $(GPU)/bt2gpu.cu

$(PKG)/libbt2host.so: $(wildcard $(HOST)/*.cpp $(HOST)/*.hpp) include/bt2gpu.h include/bt2host.h
	g++ $(CXXFLAGS) -O3 -shared -o $@ $(HOST)/host_index.cpp $(HOST)/host_reads.cpp $(HOST)/host_sink.cpp $(HOST)/host_api.cpp

oracle: $(PKG)/libbt2gpu.so
	$(MAKE) -C oracle all

emu: tests/emu/emu_pipeline tests/emu/emu_fill tests/emu/libbt2gpu_shim.so tests/emu/emu_pipeline_long tests/emu/emu_fill_long
tests/emu/emu_fill_long: tests/emu/emu_fill.cpp tests/emu/fill_scalar.hpp $(wildcard $(GPU)/*.cuh)
	g++ -O1 -std=c++17 -Iinclude $(LONGDEF) -w -o $@ tests/emu/emu_fill.cpp
tests/emu/emu_pipeline_long: tests/emu/emu_pipeline.cpp tests/emu/fill_scalar.hpp $(wildcard $(GPU)/*.cuh oracle/*.hpp $(HOST)/*.cpp)
	g++ $(CXXFLAGS) $(LONGDEF) -Ioracle -o $@ tests/emu/emu_pipeline.cpp $(HOST)/host_index.cpp $(HOST)/host_reads.cpp $(HOST)/host_sink.cpp $(HOST)/host_api.cpp
tests/emu/libbt2gpu_shim.so: tests/emu/bt2gpu_oracle_shim.cpp oracle/oracle_capi.cpp $(wildcard oracle/*.hpp) include/bt2gpu.h
	g++ $(CXXFLAGS) -Ioracle -shared -o $@ tests/emu/bt2gpu_oracle_shim.cpp oracle/oracle_capi.cpp
tests/emu/emu_fill: tests/emu/emu_fill.cpp tests/emu/fill_scalar.hpp $(wildcard $(GPU)/*.cuh)
	g++ -O1 -std=c++17 -Iinclude -w -o $@ tests/emu/emu_fill.cpp
tests/emu/emu_pipeline: tests/emu/emu_pipeline.cpp tests/emu/fill_scalar.hpp $(wildcard $(GPU)/*.cuh oracle/*.hpp $(HOST)/*.cpp)
	g++ $(CXXFLAGS) -Ioracle -o $@ tests/emu/emu_pipeline.cpp $(HOST)/host_index.cpp $(HOST)/host_reads.cpp $(HOST)/host_sink.cpp $(HOST)/host_api.cpp

clean:
	rm -f $(PKG)/*.so oracle/liboracle.so oracle/oracle_align tests/emu/emu_pipeline tests/emu/emu_fill tests/emu/emu_pipeline_long tests/emu/emu_fill_long oracle/liboracle_long.so

.PHONY: all oracle emu clean
