set -x
N=${1:-2}
timeout 600 python -m pytest tests/test_gpu_peer.py -m gpu -x -q -s 2>&1 | tail -12
for C in ${COLLS:-peer nccl}; do
  PML_BENCH_COLLECTIVE=$C timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 30 --warmup 5 --no-cpu-baseline --no-convergence 2>gpurun_out/bench_n${N}_$C.err | tail -1 > gpurun_out/bench_n${N}_$C.json
  python - <<PY
import json
j=json.loads(open("gpurun_out/bench_n${N}_$C.json").read())
print("$C", j["value"], j["ms_per_step"], j["e2e"]["value"], j.get("hessian_vector"), j["config"].get("collective"))
print(j.get("kernel_ms") or j.get("profile"))
PY
done
