# quick GPU check: parity tests + one bench line (used between kernel changes)
mkdir -p gpurun_out/q
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/q/pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/q/pytest.log
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/q/bench.json 2> gpurun_out/q/bench.err
tail -4 gpurun_out/q/pytest.log; python - <<'PY'
import json
l=json.load(open('gpurun_out/q/bench.json'))
print({k:l[k] for k in ('value','ms_per_step','stage_ms_per_step')}, l['e2e']['value'], l['roofline']['frac'], l.get('f64_rescan_cells_per_step'))
PY
