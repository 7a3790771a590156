set -x
N=${1:-2}
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 tools/bench_merge.py 2>&1 | grep -v "^\*\|OMP_NUM\|^$\|NCCL version" | tail -5
