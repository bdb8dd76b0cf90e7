#!/bin/bash
# One shard of eight (the floor of the single-frame strong scaling) with every library under build_ab/ against the default
mkdir -p gpurun_out
for lib in default $(ls build_ab/*.so 2>/dev/null) default; do
  if [ "$lib" = default ]; then unset SRT_LIB; else export SRT_LIB=$PWD/$lib; fi
  for q in 1 0; do
    export SRT_SHADOW_QUAD=$q
    echo "== $lib SRT_SHADOW_QUAD=$q"
    for a in 0 180; do python tools/shard_floor.py 8 $a | grep "shard [01] of 8"; done
  done
done 2>&1 | tee gpurun_out/ab_shard.txt
