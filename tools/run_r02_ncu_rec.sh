# recognition evidence: launch list of `bench.py --workload recognize` + one --set full capture of the search kernel and
# of the dense refinement at that shape (each after the plain command has run without ncu)
mkdir -p gpurun_out/rr
timeout 300 python bench.py --workload recognize --steps 2 --warmup 2 > gpurun_out/rr/bench.json 2> gpurun_out/rr/bench.err || exit 1
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/rr/launches.csv python bench.py --workload recognize --steps 2 --warmup 1 > gpurun_out/rr/ncu_list.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_recognize_reg -s 2 -c 1 -o gpurun_out/rr/k_recognize_reg -f python tools/rec_ab.py bench > gpurun_out/rr/ncu_rec.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tc_refine_sm -s 2 -c 1 -o gpurun_out/rr/k_tc_refine_sm_dense -f python tools/rec_ab.py bench > gpurun_out/rr/ncu_ref.log 2>&1
ls -la gpurun_out/rr
