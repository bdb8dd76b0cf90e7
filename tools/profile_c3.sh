#!/bin/bash
# ncu --set full of the three render kernels of one C3 frame at 180 degrees (one piece).  usage: tools/profile_c3.sh <tag>
T=${1:-x}
C3="python tools/profile_frame.py --config C3 --angle 180 --bands 1 --frames 3"
$C3 > gpurun_out/${T}_c3_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_phase -s 6 -c 3 -f -o gpurun_out/${T}_c3_stages $C3 > gpurun_out/${T}_c3_full.log 2>&1
cat gpurun_out/${T}_c3_plain.log | tail -1
