#!/bin/bash
# primary-pass A/B: per-pass times of C3 at 0 / 90 / 180 degrees, C2, C4 (one piece) and the two-band C3 frame at 0 degrees
for lib in default $(ls build_ab/*.so 2>/dev/null); do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for a in 0 90 180; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 5 | tail -1 | sed 's/.*frame 4: //'; done
  python tools/profile_frame.py --config C2 --angle 0 --bands 1 --frames 5 | tail -1 | sed 's/.*frame 4: /C2 /'
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 5 | tail -1 | sed 's/.*frame 4: /C4 /'
  python tools/profile_frame.py --config C3 --angle 0 --bands 2 --frames 5 | tail -1 | sed 's/.*frame 4: /banded /'
done
