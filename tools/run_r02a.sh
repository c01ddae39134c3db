# round 2, first GPU call: everything changed so far
mkdir -p gpurun_out/a
timeout 1700 python -m pytest tests -m gpu -x -q -s > gpurun_out/a/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/a/pytest.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/a/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/a/smoke.log
timeout 600 python bench.py > gpurun_out/a/bench.json 2> gpurun_out/a/bench.err; echo "bench rc=$?" >> gpurun_out/a/bench.err
timeout 300 python tools/time_mstep.py > gpurun_out/a/mstep.json 2>&1
tail -5 gpurun_out/a/pytest.log; tail -3 gpurun_out/a/smoke.log; tail -3 gpurun_out/a/bench.err; cat gpurun_out/a/mstep.json
python - <<'PY'
import json
l=json.load(open('gpurun_out/a/bench.json'))
print({k:l[k] for k in ('value','ms_per_step','stage_ms_per_step')}, l['e2e']['value'], l['roofline']['frac'], l.get('parity_check'))
PY
