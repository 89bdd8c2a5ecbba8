#!/bin/bash
# GPU session r01e: DFMA probe, GPU parity tests, bench A/B of the engine-v3 switches, launch list, ncu captures.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r01e_smi.txt
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -o /tmp/dfma_fe tools/ubench/dfma_fe.cu && timeout 300 /tmp/dfma_fe > gpurun_out/r01e_dfma_fe.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r01e_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01e_pytest.txt
tail -5 gpurun_out/r01e_pytest.txt
for B in 2048 4096; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --batch $B > gpurun_out/r01e_bench_B$B.json 2> gpurun_out/r01e_bench_B$B.err
done
BPP_NO_OVERLAP=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01e_bench_nooverlap.json 2>/dev/null
BPP_LIB=$PWD/bulletproofsplus_b200/libbpp_b200_int.so timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01e_bench_intacc.json 2>/dev/null
for c in "prove 256" "verify 2048" "verify_rlc 2048" "msm 22" "msm 16"; do
  timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01e_cases.txt 2>&1
  BPP_NO_SPLIT=1 BPP_NO_OVERLAP=1 timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01e_cases_old.txt 2>&1
  BPP_LIB=$PWD/bulletproofsplus_b200/libbpp_b200_int.so timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01e_cases_intacc.txt 2>&1
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r01e_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01e_ncu_bench.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01e_launches_prove_B256.csv python tools/run_case.py prove 256 1 > gpurun_out/r01e_ncu_prove.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"msm_sort_job|msm_reduce_chunks|msm_combine|msm_accumulate_kernel" -s 31 -c 7 -o gpurun_out/r01e_small_kernels python tools/run_case.py verify 2048 1 > gpurun_out/r01e_ncu_full.log 2>&1
ls -la gpurun_out
