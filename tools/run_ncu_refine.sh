# one ncu --set full capture of the state-major refinement at the bench shape (tools/rem_ab.py drives it) + durations of the rescan
mkdir -p gpurun_out/n
timeout 600 ncu --set full --clock-control none --import-source on -k regex:${1:-k_tc_refine_sm} -s 3 -c 1 -o gpurun_out/n/refine -f python tools/rem_ab.py > gpurun_out/n/ncu_r.log 2>&1
tail -3 gpurun_out/n/ncu_r.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"k_tc_refine|k_tc_rescan" -s 4 -c 4 --csv --log-file gpurun_out/n/refine_list.csv python tools/rem_ab.py > gpurun_out/n/ncu_r2.log 2>&1
grep -o '"void [^"]*\|"gpu__time_duration.sum","ns","[0-9]*' gpurun_out/n/refine_list.csv | paste - - | cut -c1-160
