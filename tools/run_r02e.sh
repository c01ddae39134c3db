mkdir -p gpurun_out/e
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/e/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/e/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/e/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/e/smoke.log
timeout 900 python bench.py > gpurun_out/e/bench.json 2> gpurun_out/e/bench.err; echo "bench rc=$?" >> gpurun_out/e/bench.err
tail -4 gpurun_out/e/pytest.log; tail -2 gpurun_out/e/smoke.log; tail -3 gpurun_out/e/bench.err
python - <<'PY'
import json
l=json.load(open('gpurun_out/e/bench.json'))
print({k:l[k] for k in ('value','ms_per_step','stage_ms_per_step')}, l['e2e']['value'], l['roofline']['frac'], l['parity_check']['ok'])
for k,v in l.get('other_workloads',{}).items(): print(k, v['value'], v['ms'], v['parity_check']['ok'], v['roofline']['frac'])
PY
