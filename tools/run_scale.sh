# weak-scaling bench at N = 1, 2, 4, 8 on one box (launched as the driver does) + the two-rank NCCL test
mkdir -p gpurun_out/s
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/s/bench_n1.json 2> gpurun_out/s/bench_n1.err
for n in 2 4 8; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) bench.py --gpus $n > gpurun_out/s/bench_n$n.json 2> gpurun_out/s/bench_n$n.err
done
timeout 900 python -m pytest tests/test_multigpu.py -m gpu -q > gpurun_out/s/pytest_multigpu.log 2>&1
python - <<'PY'
import json
for n in (1, 2, 4, 8):
    try:
        l = json.loads(open('gpurun_out/s/bench_n%d.json' % n).read().strip().splitlines()[-1])
        print(n, round(l['value'] / 1e6, 1), round(l['e2e']['value'] / 1e6, 1), l['ms_per_step'], l['stage_ms_per_step'].get('allreduce'))
    except Exception as e:
        print(n, 'failed', e)
PY
tail -3 gpurun_out/s/pytest_multigpu.log
