#!/bin/bash
# A/B builds of the product library with extra -D flags:  tools/build_ab.sh <name> "<flags>"  ->  build_ab/libsrt_<name>.so
# Select at run time with SRT_LIB=build_ab/libsrt_<name>.so (simple_raytracer_b200/capi.py).
set -e
cd "$(dirname "$0")/.."
mkdir -p build_ab
CSRC=simple_raytracer_b200/csrc
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -fmad=false --prec-div=true --prec-sqrt=true --ftz=false \
  -Xcompiler -fPIC,-O2,-ffp-contract=off -Iinclude -I$CSRC -I$CSRC/host $2 -shared -cudart static -o build_ab/libsrt_$1.so \
  $CSRC/srt_capi.cu $CSRC/srt_kernels.cu $CSRC/srt_phased.cu $CSRC/srt_prep.cu $CSRC/srt_prep_capi.cu $CSRC/srt_mgpu.cu $CSRC/host/*.cpp
echo built build_ab/libsrt_$1.so
