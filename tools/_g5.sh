timeout 600 python -m pytest tests/test_gpu_md.py -m gpu -x -q 2>&1 | tail -30
for g in 0 1; do
SEQM_XL_GRAPH=$g timeout 600 python bench.py --workload c60xl --steps 20 --warmup 5 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['gpu_launches'], {k:round(v['ms'],3) for k,v in d['profile_one_step'].items()})"
done
