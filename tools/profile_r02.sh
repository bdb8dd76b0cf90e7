#!/bin/bash
# Round-2 profile captures (run under gpurun, one GPU).  Outputs under gpurun_out/; summaries are made here afterwards
# with tools/ncu_summary.py and committed under profiles/.
set -x
T=${1:-base}
mkdir -p gpurun_out
C3="python tools/profile_frame.py --config C3 --angle 180 --bands 1 --frames 3"
C4="python tools/profile_frame.py --config C4 --angle 0 --bands 1 --frames 3"
$C3 > gpurun_out/${T}_c3_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 60 --csv --log-file gpurun_out/${T}_c3_launches.csv $C3 > gpurun_out/${T}_c3_launches.log 2>&1
$C3 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_phase -s 6 -c 3 -f -o gpurun_out/${T}_c3_stages $C3 > gpurun_out/${T}_c3_full.log 2>&1
$C4 > gpurun_out/${T}_c4_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none -s 0 -c 60 --csv --log-file gpurun_out/${T}_c4_launches.csv $C4 > gpurun_out/${T}_c4_launches.log 2>&1
$C4 > /dev/null 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:render_phase -s 6 -c 3 -f -o gpurun_out/${T}_c4_stages $C4 > gpurun_out/${T}_c4_full.log 2>&1
tail -n 1 gpurun_out/${T}_c3_plain.log; tail -n 1 gpurun_out/${T}_c4_plain.log
