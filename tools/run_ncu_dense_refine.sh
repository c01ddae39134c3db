#!/bin/bash
# one ncu --set full capture of the DENSE state-major refinement (recognition pass, tools/rec_ab.py bench)
mkdir -p gpurun_out/n
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tc_refine_sm -s 2 -c 1 -o gpurun_out/n/dense_refine -f python tools/rec_ab.py bench > gpurun_out/n/ncu_dr.log 2>&1
tail -3 gpurun_out/n/ncu_dr.log
