#!/bin/bash
for b in 1 2 3 4; do for a in 0 90 180 270; do python tools/profile_frame.py --config C3 --angle $a --bands $b --frames 5 | tail -1 | sed "s/.*bands=/bands=/"; done; done
for b in 1 2 3 4; do python tools/profile_frame.py --config C4 --angle 0 --bands $b --frames 5 | tail -1 | sed "s/.*bands=/C4 bands=/"; done
