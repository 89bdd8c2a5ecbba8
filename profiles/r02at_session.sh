#!/bin/bash
# GPU session r02at (--gpus 2): world-2 NCCL parity test against the oracle, then the bench at N = 2.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_multirank.py -m gpu -x -q > gpurun_out/r02at_pytest_multirank.txt 2>&1; echo "rc=$?" >> gpurun_out/r02at_pytest_multirank.txt
tail -3 gpurun_out/r02at_pytest_multirank.txt
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 > gpurun_out/r02at_bench_n2.json 2> gpurun_out/r02at_bench_n2.err; echo "bench rc=$?"
tail -3 gpurun_out/r02at_bench_n2.err; head -c 300 gpurun_out/r02at_bench_n2.json
