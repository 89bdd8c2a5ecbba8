#!/bin/bash
# GPU session r01g: decode on the auxiliary stream; full bench line; launch lists; --set full captures of the three largest kernels.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01g_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01g_pytest.txt
tail -3 gpurun_out/r01g_pytest.txt
timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01g_bench_noextras.json 2> gpurun_out/r01g_bench_noextras.err
timeout 1500 python bench.py > gpurun_out/r01g_bench_n1.json 2> gpurun_out/r01g_bench_n1.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01g_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01g_ncu_bench.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:msm_(sort_job|reduce_chunks|accumulate)_kernel' -s 6 -c 3 -o gpurun_out/r01g_verify_kernels python tools/run_case.py verify 2048 1 > gpurun_out/r01g_ncu_full.log 2>&1
ncu -i gpurun_out/r01g_verify_kernels.ncu-rep --page details > gpurun_out/r01g_verify_kernels_details.txt 2>&1
ncu -i gpurun_out/r01g_verify_kernels.ncu-rep --page raw --csv > gpurun_out/r01g_verify_kernels_raw.csv 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/r01g_smoke.txt 2>&1
ls -la gpurun_out | tail -20
