N=$1
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29571 bench.py --config 2 --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_c2_n$N.json 2> gpurun_out/r02_c2_n$N.err; tail -2 gpurun_out/r02_c2_n$N.err; wc -c gpurun_out/r02_c2_n$N.json
