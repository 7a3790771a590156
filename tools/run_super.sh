set -x
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>gpurun_out/bench_g.err | tee gpurun_out/bench_g.json | python -c "
import json,sys
j=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(j['value'], j['ms_per_step'], j['e2e']); print(json.dumps(j.get('convergence'), indent=1)); print(json.dumps(j.get('hessian_vector'), indent=1))"
python tools/run_config2_shard.py 2>&1 | tail -5
