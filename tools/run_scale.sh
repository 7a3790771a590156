# bench.py at N ranks with the peer-memory merge (one line of output); usage: bash tools/run_scale.sh N
N=${1:-2}
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-convergence 2>gpurun_out/bench_scale_n$N.err | tail -1 > gpurun_out/bench_scale_n$N.json
python - <<PY
import json
j=json.loads(open("gpurun_out/bench_scale_n$N.json").read())
print("N=$N", round(j["value"],1), "evals/s", round(j["ms_per_step"],4), "ms; e2e", round(j["e2e"]["value"],1), "; Hv", j["hessian_vector"]["cached_evals_per_s"], j["config"]["collective"][:40])
PY
