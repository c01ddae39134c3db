mkdir -p gpurun_out/c
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/c/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/c/pytest.log
timeout 900 python bench.py > gpurun_out/c/bench.json 2> gpurun_out/c/bench.err; echo "bench rc=$?" >> gpurun_out/c/bench.err
timeout 600 python bench.py --workload stress --no-cpu-baseline > gpurun_out/c/bench_stress.json 2> gpurun_out/c/bench_stress.err; echo "rc=$?" >> gpurun_out/c/bench_stress.err
tail -4 gpurun_out/c/pytest.log; tail -5 gpurun_out/c/bench.err; tail -3 gpurun_out/c/bench_stress.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/c/bench.json'))
print({k:l[k] for k in ('value','ms_per_step','stage_ms_per_step')}, l['e2e'], l['roofline']['frac'])
for k,v in l.get('other_workloads',{}).items(): print(k, {a:v[a] for a in ('value','e2e_value','ms','stage_ms','parity_check','cpu_baseline')}, v['roofline']['frac'])
PY
