#!/bin/bash
# GPU session r02ay (--gpus 8): the bench at N = 8 and N = 4.
set -x
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29552 bench.py --gpus 8 > gpurun_out/r02ay_bench_n8.json 2> gpurun_out/r02ay_bench_n8.err; echo "bench rc=$?"
tail -3 gpurun_out/r02ay_bench_n8.err; head -c 300 gpurun_out/r02ay_bench_n8.json
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29553 bench.py --gpus 4 > gpurun_out/r02ay_bench_n4.json 2> gpurun_out/r02ay_bench_n4.err; echo "bench rc=$?"
tail -3 gpurun_out/r02ay_bench_n4.err; head -c 300 gpurun_out/r02ay_bench_n4.json
