# Top-level build: the CUDA engine (libnbody_b200.so, sm_100a only) and the C host programs
# that keep the reference's command lines (Makefile:14-61 of the reference names the binaries).
PKG := n-body-simulation_b200
CSRC := $(PKG)/csrc
HOST := $(PKG)/host
NVCC ?= nvcc
CC := gcc
NVFLAGS := --cudart static -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC
LIB := $(PKG)/libnbody_b200.so
BIN := $(PKG)/bin
CU := $(CSRC)/nb_engine.cu $(CSRC)/nb_direct_f32.cu $(CSRC)/nb_direct_f64.cu $(CSRC)/nb_aux.cu $(CSRC)/nb_bh.cu $(CSRC)/nb_bh_sort.cu
HDR := $(wildcard $(CSRC)/*.cuh) include/nbody_engine.h

.PHONY: all lib host clean
all: lib host

lib: $(LIB)
$(LIB): $(CU) $(HDR)
	$(NVCC) $(NVFLAGS) -shared -o $@ $(CU) -ldl

host: lib
	@if [ -f $(HOST)/Makefile ]; then $(MAKE) -s -C $(HOST); fi

clean:
	rm -f $(LIB); rm -rf $(BIN)
