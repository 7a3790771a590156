timeout 1000 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-convergence 2>/dev/null | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(j['value'], j['ms_per_step'], j['e2e']['value']); print({k:round(v['ms_per_launch'],4) for k,v in j['kernels'].items() if v['launches']}); print(j['hessian_vector'])"
