#!/bin/bash
# GPU session r01k (final state of round 1): tests, smoke, full bench line, launch list of the bench command.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01k_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01k_pytest.txt
tail -3 gpurun_out/r01k_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r01k_smoke.txt 2>&1
timeout 1500 python bench.py > gpurun_out/r01k_bench_n1.json 2> gpurun_out/r01k_bench_n1.err
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r01k_bench_reference.json 2> gpurun_out/r01k_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01k_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01k_ncu_bench.log 2>&1
ls -la gpurun_out | tail -8
