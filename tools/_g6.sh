timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r02b_bench_n1.json 2> gpurun_out/r02b_bench_n1.err; tail -3 gpurun_out/r02b_bench_n1.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02b_bench_n1.json').read().strip().splitlines()[-1])
print('benzene', d['value'], d['ms_per_step'], d['e2e']['value'], 'cold', d['cold_call_ms'], 'cpu', d['cpu_baseline'], 'launches', d['gpu_launches'])
x = d.get('xlbomd', {})
print('xlbomd', x.get('value'), x.get('ms_per_step'), x.get('e2e'), x.get('gpu_launches'), x.get('cpu_baseline'), x.get('roofline'))
PY
SEQM_BENCH_SUB=all SEQM_BENCH_ORGANICS=16384 timeout 900 python bench.py --steps 5 --warmup 2 > gpurun_out/r02b_bench_all.json 2> gpurun_out/r02b_bench_all.err; tail -3 gpurun_out/r02b_bench_all.err
python - <<'PY'
import json
d = json.loads(open('gpurun_out/r02b_bench_all.json').read().strip().splitlines()[-1])
for k in ('organics', 'water'):
    x = d.get(k, {})
    print(k, x.get('value'), x.get('ms_per_step'), x.get('parity'), {a: round(b['ms'], 2) for a, b in x.get('profile_one_step', {}).items()})
PY
