#!/bin/bash
# one ncu --set full capture of the state-major refinement at the stress shape (4501 x 64 densities, d = 48)
mkdir -p gpurun_out/n
timeout 700 ncu --set full --clock-control none --import-source on -k regex:k_tc_refine_sm -s 2 -c 1 -o gpurun_out/n/stress_refine -f python bench.py --workload stress --steps 1 --warmup 3 > gpurun_out/n/ncu_sr.log 2>&1
tail -3 gpurun_out/n/ncu_sr.log | cut -c1-300
