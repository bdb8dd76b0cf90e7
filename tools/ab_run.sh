#!/bin/bash
# Per-pass times of C3 at 0 / 180 degrees and C4 for the default library and every build_ab/ variant.
for lib in default $(ls build_ab/*.so 2>/dev/null); do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for a in 0 180; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 4 | tail -1 | sed 's/.*frame 3: //'; done
  python tools/profile_frame.py --config C3 --angle 180 --bands 2 --frames 4 | tail -1 | sed 's/.*frame 3: //'
done
