N=$1
T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $T --master-port 29531 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; tail -2 gpurun_out/r02_bench_n$N.err
timeout 400 $T --master-port 29532 bench.py --config 2 --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_c2_n$N.json 2> gpurun_out/r02_c2_n$N.err; tail -2 gpurun_out/r02_c2_n$N.err
timeout 600 $T --master-port 29533 bench.py --config 4 --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_c4_n$N.json 2> gpurun_out/r02_c4_n$N.err; tail -2 gpurun_out/r02_c4_n$N.err
wc -c gpurun_out/r02_*_n$N.json
