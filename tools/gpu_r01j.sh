#!/bin/bash
# GPU session r01j: sort kernel with compile-time window width (scalar in registers); tests, A/B of the two register allocations,
# full bench, launch list, sort capture.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01j_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01j_pytest.txt
tail -3 gpurun_out/r01j_pytest.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01j_bench_noextras.json 2> gpurun_out/r01j_bench_noextras.err
BPP_SORT_BLOCKS=2 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01j_bench_noextras_sort2.json 2>/dev/null
timeout 1500 python bench.py > gpurun_out/r01j_bench_n1.json 2> gpurun_out/r01j_bench_n1.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01j_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01j_ncu_bench.log 2>&1
BPP_SORT_BLOCKS=2 timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01j_launches_bench_steps2_sort2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01j_ncu_bench2.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:msm_sort_job_kernel' -s 2 -c 1 -o gpurun_out/r01j_sort python tools/run_case.py verify 2048 1 > gpurun_out/r01j_ncu_full.log 2>&1
ncu -i gpurun_out/r01j_sort.ncu-rep --page details > gpurun_out/r01j_sort_details.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/r01j_smoke.txt 2>&1
ls -la gpurun_out | tail -8
