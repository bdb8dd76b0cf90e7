#!/bin/bash
# A/B of the four-wide shadow traversal (SRT_SHADOW_QUAD=0 / 1, one library): per-pass times, then the GPU test suite with it on
mkdir -p gpurun_out
for q in 0 1 0 1; do
  export SRT_SHADOW_QUAD=$q
  echo "== SRT_SHADOW_QUAD=$q"
  for a in 0 90 180 270; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: /C4 /'
  python tools/profile_frame.py --config C3 --angle 180 --bands 2 --frames 8 | tail -1 | sed 's/.*frame 7: /banded180 /'
done 2>&1 | tee gpurun_out/ab_quad.txt
unset SRT_SHADOW_QUAD
( timeout 900 python -m pytest tests/ -x -q -m gpu 2>&1 | tail -8 ) | tee gpurun_out/ab_quad_pytest.txt
