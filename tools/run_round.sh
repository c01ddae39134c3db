set -x
mkdir -p gpurun_out/w
nvidia-smi --query-gpu=name,clocks.max.sm --format=csv > gpurun_out/w/smi.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/w/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/w/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/w/smoke.log 2>&1
timeout 600 python bench.py --impl reference > gpurun_out/w/bench_ref.json 2> gpurun_out/w/bench_ref.err
timeout 600 python bench.py > gpurun_out/w/bench.json 2> gpurun_out/w/bench.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/w/launches.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/w/ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_gemm256 -s 2 -c 1 -o gpurun_out/w/gemm -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/w/ncu_g.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_tc_refine_sm -s 1 -c 1 -o gpurun_out/w/refine -f python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/w/ncu_v.log 2>&1
tail -3 gpurun_out/w/pytest.log; cat gpurun_out/w/bench.json
