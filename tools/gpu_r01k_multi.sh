#!/bin/bash
# usage: gpu_r01k_multi.sh N   (under gpurun --gpus N)
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 20 --warmup 3 \
  > gpurun_out/r01k_bench_n$N.json 2> gpurun_out/r01k_bench_n$N.err
tail -c 600 gpurun_out/r01k_bench_n$N.json
