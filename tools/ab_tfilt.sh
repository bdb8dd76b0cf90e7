#!/bin/bash
# A/B of the shadow pass's per-triangle certain-miss records (SRT_SHADOW_TRIFILTER=0 / 1), default library (64 registers) and
# build_ab/libsrt_lb9.so (the shadow kernel held at 9 CTAs per SM = 56 registers)
mkdir -p gpurun_out
for lib in default build_ab/libsrt_lb9.so; do
for q in 0 1 0 1; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  export SRT_SHADOW_TRIFILTER=$q
  echo "== $lib SRT_SHADOW_TRIFILTER=$q"
  for a in 0 90 180 270; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: /C4 /'
  python tools/profile_frame.py --config C3 --angle 180 --bands 2 --frames 8 | tail -1 | sed 's/.*frame 7: /banded180 /'
done; done 2>&1 | tee gpurun_out/ab_tfilt.txt
unset SRT_SHADOW_TRIFILTER; unset SRT_LIB
( timeout 900 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -8 ) | tee gpurun_out/ab_tfilt_pytest.txt
