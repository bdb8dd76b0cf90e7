# Top-level build: the product library (CUDA, sm_100a only) and the oracle (tests only).
#   make            -> simple_raytracer_b200/lib/libsrt_b200.so
#   make oracle     -> oracle/liboracle.so
#   make ref        -> oracle/_ref/libsrt_ref.so (+ staged assets), needs /root/reference
NVCC     ?= /usr/local/cuda/bin/nvcc
ARCH     := -gencode arch=compute_100a,code=sm_100a
# -fmad=false / --prec-div / --prec-sqrt: belt and braces; the device code already spells every
# operation with __fmul_rn/__fadd_rn/__fdiv_rn/__fsqrt_rn.
NVFLAGS  := $(ARCH) -O3 -std=c++17 -lineinfo -fmad=false --prec-div=true --prec-sqrt=true --ftz=false \
            -Xcompiler -fPIC,-O2,-ffp-contract=off -Iinclude -Isimple_raytracer_b200/csrc -Isimple_raytracer_b200/csrc/host
CSRC     := simple_raytracer_b200/csrc
LIB      := simple_raytracer_b200/lib/libsrt_b200.so
SRCS     := $(CSRC)/srt_capi.cu $(CSRC)/srt_kernels.cu $(CSRC)/srt_phased.cu $(CSRC)/srt_prep.cu $(CSRC)/srt_prep_capi.cu $(CSRC)/srt_mgpu.cu $(wildcard $(CSRC)/host/*.cpp)
HDRS     := include/srt.h include/srt_host.h $(wildcard $(CSRC)/*.cuh) $(wildcard $(CSRC)/*.hpp) $(wildcard $(CSRC)/host/*.hpp)

.PHONY: all lib oracle ref clean
all: lib
lib: $(LIB)
$(LIB): $(SRCS) $(HDRS)
	@mkdir -p $(dir $(LIB))
	$(NVCC) $(NVFLAGS) -shared -cudart static -o $@ $(SRCS)
oracle:
	$(MAKE) -C oracle oracle
ref: lib
	$(MAKE) -C oracle ref
clean:
	rm -f $(LIB)
	$(MAKE) -C oracle clean
