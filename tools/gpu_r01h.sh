#!/bin/bash
# GPU session r01h: sort kernel with four shared atomics in flight; tests, bench, launch list, sort capture.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01h_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01h_pytest.txt
tail -3 gpurun_out/r01h_pytest.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01h_bench_noextras.json 2> gpurun_out/r01h_bench_noextras.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01h_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01h_ncu_bench.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:msm_sort_job_kernel' -s 2 -c 1 -o gpurun_out/r01h_sort python tools/run_case.py verify 2048 1 > gpurun_out/r01h_ncu_full.log 2>&1
ncu -i gpurun_out/r01h_sort.ncu-rep --page details > gpurun_out/r01h_sort_details.txt 2>&1
for c in "prove 256" "verify 2048" "msm 22" "msm 16"; do timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01h_cases.txt 2>&1; done
ls -la gpurun_out | tail -8
