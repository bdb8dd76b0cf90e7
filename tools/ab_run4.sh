#!/bin/bash
# A/B of the closest-hit triangle-test forms (build_ab/libsrt_mt1.so = filtered inline, mt2 = classification + deferred tail)
mkdir -p gpurun_out
for lib in default build_ab/libsrt_mt1.so build_ab/libsrt_mt2.so; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for rep in 1 2; do
  for a in 0 90 180 270; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 6 | tail -1 | sed 's/.*frame 5: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 6 | tail -1 | sed 's/.*frame 5: /C4 /'
  python tools/profile_frame.py --config C2 --angle 0 --bands 1 --frames 6 | tail -1 | sed 's/.*frame 5: /C2 /'
  python tools/profile_frame.py --config C3 --angle 180 --bands 2 --frames 6 | tail -1 | sed 's/.*frame 5: /banded180 /'
  done
done 2>&1 | tee gpurun_out/ab_mt.txt
for lib in build_ab/libsrt_mt2.so build_ab/libsrt_mt1.so; do
  export SRT_LIB=$PWD/$lib
  echo "== parity $lib"
  timeout 600 python -m pytest tests/test_gpu_golden.py tests/test_gpu_leaf.py tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -5
done 2>&1 | tee gpurun_out/ab_mt_parity.txt
