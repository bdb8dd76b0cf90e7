#!/bin/bash
for lib in default $(ls build_ab/*.so 2>/dev/null); do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  echo "== $lib"
  for a in 0 180; do python tools/profile_frame.py --config C3 --angle $a --bands 1 --frames 5 | tail -1 | sed 's/.*frame 4: //'; done
  python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 5 | tail -1 | sed 's/.*frame 4: /C4 /'
  python tools/profile_frame.py --config C3 --angle 0 --bands 2 --frames 5 | tail -1 | sed 's/.*frame 4: /banded /'
  python tools/shard_floor.py 8 0 | grep "shard 0 of 8"
done
