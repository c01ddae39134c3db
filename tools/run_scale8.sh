# the 8-GPU line alone (launched as the driver does) + the two-rank NCCL test
mkdir -p gpurun_out/s8
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29508 bench.py --gpus 8 > gpurun_out/s8/bench_n8.json 2> gpurun_out/s8/bench_n8.err
timeout 600 python -m pytest tests/test_multigpu.py -m gpu -q > gpurun_out/s8/pytest_multigpu.log 2>&1
python - <<'PY'
import json
l = json.loads(open('gpurun_out/s8/bench_n8.json').read().strip().splitlines()[-1])
print(round(l['value'] / 1e6, 1), round(l['e2e']['value'] / 1e6, 1), l['ms_per_step'], {k: (round(v['value'] / 1e6, 1), v['parity_check']['ok']) for k, v in l['other_workloads'].items()})
PY
tail -2 gpurun_out/s8/pytest_multigpu.log; tail -c 300 gpurun_out/s8/bench_n8.err
