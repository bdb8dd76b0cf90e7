#!/bin/bash
# Reproduces order-dependent failures: the whole -m gpu suite, N times, with two row bands as the default.
# usage: tools/flaky_hunt.sh [runs] [bands]
N=${1:-12}; B=${2:-2}
mkdir -p gpurun_out
fails=0
for i in $(seq 1 $N); do
  SRT_DEFAULT_BANDS=$B timeout 600 python -m pytest tests -m gpu -q -p no:randomly -p no:cacheprovider > gpurun_out/hunt_$i.log 2>&1
  rc=$?
  tail -1 gpurun_out/hunt_$i.log
  if [ $rc -ne 0 ]; then fails=$((fails+1)); grep -n "AssertionError\|^FAILED\|^E " gpurun_out/hunt_$i.log | head -20; else rm -f gpurun_out/hunt_$i.log; fi
done
echo "flaky_hunt: $fails failing runs of $N (bands=$B)"
