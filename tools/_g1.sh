set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -5
python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench_base.json 2> gpurun_out/r02a_bench_base.err
for tol in 1e-16 1e-14 1e-12; do
  SEQM_EIGH_TOL=$tol python bench.py --steps 20 --warmup 3 > gpurun_out/r02a_bench_tol$tol.json 2>/dev/null
  SEQM_EIGH_TOL=$tol python -m pytest tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -3 > gpurun_out/r02a_parity_tol$tol.txt
done
SEQM_BENCH_ORGANICS=16384 python bench.py --workload organics --steps 3 --warmup 2 > gpurun_out/r02a_org16k_base.json 2>/dev/null
SEQM_EIGH_TOL=1e-14 SEQM_BENCH_ORGANICS=16384 python bench.py --workload organics --steps 3 --warmup 2 > gpurun_out/r02a_org16k_tol1e-14.json 2>/dev/null
for f in gpurun_out/r02a_*.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
    print(sys.argv[1], d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('scf_iterations'), d['config'].get('scf_not_converged'))
except Exception as e: print(sys.argv[1], 'ERR', e)
PY
done
cat gpurun_out/r02a_parity_tol*.txt
