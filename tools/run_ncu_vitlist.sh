# ncu launch list of the Viterbi class launches (serialised, cold) at the bench shape
mkdir -p gpurun_out/v
timeout 600 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size,launch__shared_mem_per_block_dynamic --clock-control none -k regex:k_viterbi -c 40 --csv --log-file gpurun_out/v/vit2.csv python tools/vit_ab.py 4096 ${1:-1} > gpurun_out/v/log2.txt 2>&1
tail -2 gpurun_out/v/log2.txt
