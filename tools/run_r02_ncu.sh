# round-2 evidence: launch list of the bench command + one --set full capture each of the filter, the refinement, the rescan and a Viterbi class
mkdir -p gpurun_out/r
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r/launches.csv python bench.py --steps 2 --warmup 1 --no-other-workloads > gpurun_out/r/ncu_list.log 2>&1
for k in k_tc_gemm256 k_tc_refine_sm k_tc_rescan2; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 1 -o gpurun_out/r/$k -f python tools/rem_ab.py > gpurun_out/r/ncu_$k.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_viterbi -s 16 -c 1 -o gpurun_out/r/k_viterbi -f python tools/vit_ab.py 4096 1 > gpurun_out/r/ncu_vit.log 2>&1
ls -la gpurun_out/r
