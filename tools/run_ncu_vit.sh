# one ncu --set full capture of a Viterbi class launch at the bench shape (tools/vit_ab.py, single-warp kernel)
mkdir -p gpurun_out/n
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_viterbi -s ${1:-5} -c 1 -o gpurun_out/n/vit -f python tools/vit_ab.py 4096 0 > gpurun_out/n/ncu_v.log 2>&1
tail -3 gpurun_out/n/ncu_v.log
