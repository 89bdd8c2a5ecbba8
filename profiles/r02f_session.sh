#!/bin/bash
# GPU session r02f: direct kernel with slices + balanced leftover. Parity subset, then step timings at three table budgets.
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_direct.py tests/test_gpu_cpp_host.py tests/test_gpu_hashed_pk.py tests/test_gpu_session.py -m gpu -x -q > gpurun_out/r02f_pytest.txt 2>&1; echo "rc=$?" >> gpurun_out/r02f_pytest.txt
tail -4 gpurun_out/r02f_pytest.txt
for gb in 48 80 130; do
  BPP_TABLE_GB=$gb timeout 600 python tools/run_case.py verify 2048 5 > gpurun_out/r02f_verify_tables_${gb}gb.txt 2>&1; tail -2 gpurun_out/r02f_verify_tables_${gb}gb.txt
done
BPP_TABLE_GB=48 timeout 600 python tools/run_case.py prove 256 5 > gpurun_out/r02f_prove_tables_48gb.txt 2>&1; tail -2 gpurun_out/r02f_prove_tables_48gb.txt
BPP_TABLE_GB=48 timeout 600 python tools/trace_case.py verify 2048 > gpurun_out/r02f_trace_verify_2048_tables.txt 2>&1
BPP_TABLE_GB=48 timeout 600 python tools/trace_case.py prove 256 > gpurun_out/r02f_trace_prove_256_tables.txt 2>&1
timeout 600 python bench.py --no-extras > gpurun_out/r02f_bench_noextras_48.json 2>/dev/null
timeout 600 python bench.py --no-extras --table-gb 80 > gpurun_out/r02f_bench_noextras_80.json 2>/dev/null
for f in gpurun_out/r02f_bench_noextras_*.json; do python -c "
import json; d=json.loads(open('$f').read()); print('$f', round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3), d['config']['fixed_base_tables']['window_bits'], round(d['config']['fixed_base_tables']['build_s'],3))"; done
