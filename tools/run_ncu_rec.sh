#!/bin/bash
# one --set full capture of the search kernel on the lone-warp case (148 utterances of 1399 frames)
set -e
timeout 200 python tools/rec_ab.py all1399_148
timeout 500 ncu --set full --clock-control none --import-source on -k regex:k_recognize -s 3 -c 1 \
  -o gpurun_out/rec_lone -f python tools/rec_ab.py all1399_148 > gpurun_out/rec_lone_ncu.log 2>&1
tail -3 gpurun_out/rec_lone_ncu.log
