# Build: product library (CUDA, sm_100a) + oracle (CPU checker).
NVCC ?= /usr/local/cuda/bin/nvcc
PKG := x-edr-pointcloud-segmentation_b200
CSRC := $(PKG)/csrc
NVFLAGS := -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 \
           -fmad=false -prec-div=true -prec-sqrt=true -ftz=false \
           -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math -shared

all: lib oracle host_test

lib: $(PKG)/libpcseg_b200.so

$(PKG)/libpcseg_b200.so: $(CSRC)/pcseg_api.cu $(wildcard $(CSRC)/*.cuh) include/pcseg_b200.h
	$(NVCC) $(NVFLAGS) $(EXTRA) -o $@.tmp $(CSRC)/pcseg_api.cu && mv -f $@.tmp $@

oracle:
	$(MAKE) -C oracle all

# C++ host-side mirror of the reference API (host/pcseg_host.h) + its parity test binary
host_test: tests/host/clustering_host_test
tests/host/clustering_host_test: tests/host/clustering_host_test.cc $(PKG)/host/pcseg_host.h include/pcseg_b200.h $(PKG)/libpcseg_b200.so
	$(CXX) -std=c++17 -O2 -Wall -Iinclude -I$(PKG)/host -o $@ $< -L$(PKG) -lpcseg_b200 -Wl,-rpath,'$$ORIGIN/../../$(PKG)'

# diagnostics build with the per-role cycle counters (tools/probe_prof.py loads it through PCSEG_LIB)
prof: build/libpcseg_prof.so
build/libpcseg_prof.so: $(CSRC)/pcseg_api.cu $(wildcard $(CSRC)/*.cuh) include/pcseg_b200.h
	mkdir -p build && $(NVCC) $(NVFLAGS) -DPCSEG_PROFILE -o $@ $(CSRC)/pcseg_api.cu

clean:
	rm -f $(PKG)/libpcseg_b200.so; $(MAKE) -C oracle clean
.PHONY: all lib oracle clean host_test prof
