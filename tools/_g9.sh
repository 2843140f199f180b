timeout 900 python -m pytest tests -m gpu -q -x 2>&1 | tail -6
SEQM_BENCH_SUB=0 python bench.py --steps 20 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('benzene', d['value'], d['ms_per_step'], d['e2e']['value'])"
SEQM_BENCH_ORGANICS=16384 python bench.py --workload organics --steps 3 --warmup 2 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('organics16k', d['value'], d['ms_per_step'], {k: round(v['ms'],1) for k,v in d['profile_one_step'].items() if v['ms']>2})"
