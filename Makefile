# Builds every native piece in-tree: the CUDA product library, the C++ host library, the CPU oracle.
NVCC ?= /usr/local/cuda/bin/nvcc
CXX ?= g++
CSRC := raytracer2_b200/csrc
CU_SRCS := $(CSRC)/rt2_api.cu $(CSRC)/rt2_scene.cu $(CSRC)/rt2_build.cu $(CSRC)/rt2_render.cu $(CSRC)/rt2_noalloc.cu $(CSRC)/rt2_group.cu $(wildcard $(CSRC)/rt2_passes*.cu) $(wildcard $(CSRC)/rt2_pass_*.cu)
CU_HDRS := $(wildcard $(CSRC)/*.cuh) include/rt2_abi.h include/rt2_gconst.h
# -fmad=false: no implicit FMA contraction (bit parity with the oracle); explicit __fmaf_rn where wanted.
NVCCFLAGS := $(EXTRA_NVCCFLAGS) -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false -Xcompiler -fPIC,-ffp-contract=off,-Wall -Xptxas -v

all: raytracer2_b200/librt2cuda.so raytracer2_b200/librt2host.so oracle

# one object per translation unit (no relocatable device code: no kernel calls across files), so `make -j` builds in parallel and
# an edit recompiles one file; every unit's ptxas -v output is kept in build/<unit>.ptxas.log and joined into build_ptxas.log
OBJDIR := build/$(if $(EXTRA_NVCCFLAGS),variant,default)
CU_OBJS := $(patsubst $(CSRC)/%.cu,$(OBJDIR)/%.o,$(CU_SRCS))
$(OBJDIR)/%.o: $(CSRC)/%.cu $(CU_HDRS)
	@mkdir -p $(OBJDIR)
	$(NVCC) $(NVCCFLAGS) -Wno-deprecated-gpu-targets -Xcompiler -Wno-unknown-pragmas -c -o $@ $< 2> $(OBJDIR)/$*.ptxas.log || (cat $(OBJDIR)/$*.ptxas.log; exit 1)
	@grep -E "error|warning" $(OBJDIR)/$*.ptxas.log || true

raytracer2_b200/librt2cuda.so: $(CU_OBJS)
	$(NVCC) -shared -o $@ $(CU_OBJS) $(LDLIBS_CUDA) -ldl
	@cat $(sort $(wildcard $(OBJDIR)/*.ptxas.log)) > build_ptxas.log

raytracer2_b200/librt2host.so: raytracer2_b200/host/rt2_host.cpp raytracer2_b200/host/rt2_ingest.cpp raytracer2_b200/host/rt2_ingest_codecs.cpp raytracer2_b200/host/rt2_host.h raytracer2_b200/host/rt2_ingest.h include/rt2_abi.h include/rt2_gconst.h
	$(CXX) -O2 -std=c++17 -fPIC -Wall -Wextra -shared -o $@ raytracer2_b200/host/rt2_host.cpp raytracer2_b200/host/rt2_ingest.cpp raytracer2_b200/host/rt2_ingest_codecs.cpp -lz

oracle:
	$(MAKE) -C oracle

clean:
	rm -rf raytracer2_b200/librt2cuda.so raytracer2_b200/librt2host.so build_ptxas.log build
	$(MAKE) -C oracle clean

.PHONY: all oracle clean
