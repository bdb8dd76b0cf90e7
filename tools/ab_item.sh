#!/bin/bash
# A/B of every library under build_ab/ against the default (per-pass times, two-band frame), then the parity tests with the first of them
mkdir -p gpurun_out
for lib in default $(ls build_ab/*.so 2>/dev/null) default; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for rep in 1 2; do
  for a in 0 90 180 270; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: /C4 /'
  python tools/profile_frame.py --config C3 --angle 180 --bands 2 --frames 8 | tail -1 | sed 's/.*frame 7: /banded180 /'
  done
done 2>&1 | tee gpurun_out/ab_item.txt
export SRT_LIB=$PWD/$(ls build_ab/*.so | head -1)
( timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_edges.py tests/test_gpu_mgpu.py -x -q -m gpu 2>&1 | tail -5 ) | tee gpurun_out/ab_item_pytest.txt
