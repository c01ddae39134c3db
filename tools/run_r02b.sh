mkdir -p gpurun_out/b
timeout 1700 python -m pytest tests -m gpu -x -q -s > gpurun_out/b/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/b/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/b/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/b/smoke.log
timeout 600 python bench.py > gpurun_out/b/bench.json 2> gpurun_out/b/bench.err; echo "bench rc=$?" >> gpurun_out/b/bench.err
timeout 300 python tools/time_mstep.py > gpurun_out/b/mstep.json 2>&1
grep -v "^digits\|^\.digits" gpurun_out/b/pytest.log | tail -25; tail -3 gpurun_out/b/smoke.log; tail -5 gpurun_out/b/bench.err; cat gpurun_out/b/mstep.json
python - <<'PY'
import json
l=json.load(open('gpurun_out/b/bench.json'))
print({k:l[k] for k in ('value','ms_per_step','stage_ms_per_step')}, l['e2e'], l['roofline']['frac'], l.get('parity_check'))
PY
