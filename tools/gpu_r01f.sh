#!/bin/bash
# GPU session r01f: engine v3 on the integer pipe (folding, split digits, priority overlap), A/B switches, batch sizes, ncu.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r01f_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r01f_pytest.txt
tail -3 gpurun_out/r01f_pytest.txt
for B in 2048 4096 8192; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-extras --batch $B > gpurun_out/r01f_bench_B$B.json 2> gpurun_out/r01f_bench_B$B.err
done
BPP_NO_OVERLAP=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01f_bench_nooverlap.json 2>/dev/null
BPP_HYBRID=5 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01f_bench_hybrid5.json 2>/dev/null
BPP_HYBRID=1 timeout 600 python bench.py --steps 10 --warmup 3 --no-extras > gpurun_out/r01f_bench_hybrid1.json 2>/dev/null
for c in "prove 256" "verify 2048" "verify_rlc 2048" "msm 22" "msm 16" "prove_fs 256"; do
  timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01f_cases.txt 2>&1
  BPP_NO_SPLIT=1 BPP_NO_OVERLAP=1 timeout 300 python tools/run_case.py $c 5 >> gpurun_out/r01f_cases_nosplit_nooverlap.txt 2>&1
done
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r01f_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r01f_ncu_bench.log 2>&1
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r01f_launches_prove_B256.csv python tools/run_case.py prove 256 1 > gpurun_out/r01f_ncu_prove.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:.*msm_(sort_job|reduce_chunks|accumulate)_kernel.*' -s 27 -c 6 -o gpurun_out/r01f_verify_kernels python tools/run_case.py verify 2048 1 > gpurun_out/r01f_ncu_full.log 2>&1
ls -la gpurun_out | tail -20
