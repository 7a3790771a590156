N=$1
if [ "$N" = "1" ]; then T="python"; else T="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29581"; fi
timeout 900 $T bench.py --config 4 --gpus $N --steps 10 --warmup 3 > gpurun_out/r02_c4_n$N.json 2> gpurun_out/r02_c4_n$N.err; tail -2 gpurun_out/r02_c4_n$N.err
wc -c gpurun_out/r02_c4_n$N.json
