#!/bin/bash
# One ncu --set full capture of the three render kernels of the C3 180-degree frame (one piece), for the L1 / LSU side of the roofline.
mkdir -p gpurun_out
C3="python tools/profile_frame.py --config C3 --angle 180 --bands 1 --frames 3"
$C3 > gpurun_out/r02f_c3_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_phase -s 6 -c 3 -f -o gpurun_out/r02f_c3_stages $C3 > gpurun_out/r02f_c3_full.log 2>&1
tail -n 1 gpurun_out/r02f_c3_plain.log; ls -la gpurun_out/*.ncu-rep
