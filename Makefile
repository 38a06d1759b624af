# Builds the product library (nodexa_chain_core_b200/libnxkawpow.so) for sm_100a and the CPU oracle.
# nvcc cross-compiles without a GPU.  `python -c "import __graft_entry__ as g; g.build()"` drives this.
NVCC      ?= /usr/local/cuda/bin/nvcc
CUDA_HOME ?= /usr/local/cuda
PKG       := nodexa_chain_core_b200
SRC       := $(PKG)/csrc
OBJ       := build/obj
LIB       := $(PKG)/libnxkawpow.so
ARCH      := -gencode arch=compute_100a,code=sm_100a
NVFLAGS   := $(ARCH) -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-Wall,-Wno-unused-function,-Wno-unknown-pragmas -Iinclude -I$(SRC)
CXXSRC    := nx_host.cpp nx_codegen.cpp nx_kawpow_rt.cpp nx_abi.cpp nx_kernel_src.cpp nx_tools.cpp nx_compat.cpp nx_miner.cpp
CUSRC     := nx_dag.cu nx_kawpow_interp.cu nx_equihash.cu
OBJS      := $(addprefix $(OBJ)/,$(CXXSRC:.cpp=.o) $(CUSRC:.cu=.o)) $(OBJ)/nx_lightcache_avx512.o $(OBJ)/nx_cpu_dispatch.o

.PHONY: all lib oracle ref clean ubench
all: lib oracle

lib: $(LIB)

$(SRC)/kawpow_kernel_src.inc: $(SRC)/kawpow_kernel.cuh
	( printf 'R"NXSRC(' ; cat $< ; printf ')NXSRC"\n' ) > $@

$(OBJ)/nx_kernel_src.o: $(SRC)/kawpow_kernel_src.inc
$(OBJ)/nx_kawpow_interp.o: $(SRC)/kawpow_kernel.cuh
$(OBJ)/nx_equihash.o: $(SRC)/nx_blake2b.cuh include/nx_equihash.h

# plain host code built by g++ directly: the AVX-512 light-cache chain (only called when the CPU has AVX-512F)
$(OBJ)/nx_lightcache_avx512.o: $(SRC)/nx_lightcache_avx512.cpp
	@mkdir -p $(OBJ)
	g++ -O3 -std=c++17 -fPIC -Wall -mavx512f -c $< -o $@
$(OBJ)/nx_cpu_dispatch.o: $(SRC)/nx_cpu_dispatch.cpp
	@mkdir -p $(OBJ)
	g++ -O2 -std=c++17 -fPIC -Wall -c $< -o $@

$(OBJ)/%.o: $(SRC)/%.cpp $(SRC)/nx_internal.hpp $(SRC)/nx_keccak.cuh include/nx_kawpow.h include/nx/ethash_compat.h include/nx/ethash_compat.hpp
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -x cu -c $< -o $@

$(OBJ)/%.o: $(SRC)/%.cu $(SRC)/nx_internal.hpp $(SRC)/nx_keccak.cuh include/nx_kawpow.h include/nx/ethash_compat.h include/nx/ethash_compat.hpp
	@mkdir -p $(OBJ)
	$(NVCC) $(NVFLAGS) -Xptxas -v -c $< -o $@

$(LIB): $(OBJS)
	$(NVCC) $(ARCH) -shared -cudart static -o $@ $(OBJS) -L$(CUDA_HOME)/lib64 -lnvrtc -ldl -lpthread \
	    -Xlinker -rpath=$(CUDA_HOME)/lib64

oracle:
	$(MAKE) -C oracle oracle

# stand-alone micro-benchmarks (tools/ubench): the measured ceilings DESIGN.md section 4 divides by
ubench: build/nx_ubench
build/nx_ubench: tools/ubench/nx_ubench.cu
	@mkdir -p build
	$(NVCC) $(ARCH) -O3 -lineinfo -std=c++17 -cudart static -o $@ $<

ref: lib
	$(MAKE) -C oracle ref ref_tests ref_node_dropin ref_node_explicit

clean:
	rm -rf build $(LIB) $(SRC)/kawpow_kernel_src.inc
	$(MAKE) -C oracle clean
