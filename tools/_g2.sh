python -m pytest tests/test_gpu_golden_big.py -m gpu -q -x --deselect tests/test_gpu_golden_big.py::test_eigh_above_118_matches_reference 2>&1 | tail -40
python -m pytest tests -m gpu -q --deselect tests/test_gpu_golden_big.py 2>&1 | tail -8
python bench.py --steps 20 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
python bench.py --workload c60xl --steps 20 --warmup 3 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:v['ms'] for k,v in d['profile_one_step'].items()})"
python bench.py --workload water --steps 3 --warmup 2 | python -c "import json,sys; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], {k:v['ms'] for k,v in d['profile_one_step'].items()})"
