# one ncu --set full capture of the tensor-core filter at the bench shape (tools/rem_ab.py drives it: cheap set-up)
mkdir -p gpurun_out/n
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_tc_gemm256 -s 6 -c 1 -o gpurun_out/n/gemm -f python tools/rem_ab.py > gpurun_out/n/ncu_g.log 2>&1
tail -3 gpurun_out/n/ncu_g.log
