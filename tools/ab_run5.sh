#!/bin/bash
# A/B of every library under build_ab/ against the default: per-pass times of C3 at 0 / 180 degrees and of C4 (one piece), two repetitions
mkdir -p gpurun_out
for lib in default $(ls build_ab/*.so 2>/dev/null) default; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for rep in 1 2; do
  for a in 0 180; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 8 | tail -1 | sed 's/.*frame 7: /C4 /'
  done
done 2>&1 | tee gpurun_out/ab5.txt
