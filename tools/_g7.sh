timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r02b_bench_n2.json 2> gpurun_out/r02b_bench_n2.err; tail -5 gpurun_out/r02b_bench_n2.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02b_bench_n2.json').read().strip().splitlines()[-1])
print('benzene', d['value'], d['ms_per_step'], d['e2e']['value'])
for k in ('xlbomd', 'organics', 'water'):
    x = d.get(k, {})
    print(k, x.get('value'), x.get('ms_per_step'), x.get('e2e', {}).get('value'), x.get('parity'))
PY
timeout 600 python -m pytest tests/test_gpu_sp2_sharded.py -m gpu -q 2>&1 | tail -5
