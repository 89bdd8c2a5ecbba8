#!/bin/bash
# r02n: scalar multiply / square as real functions (instruction-fetch-bound scalar kernels)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_session.py tests/test_gpu_direct.py -m gpu -x -q > gpurun_out/r02n_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02n_pytest.txt
tail -2 gpurun_out/r02n_pytest.txt
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify_fs 2048 > gpurun_out/r02n_trace_verify_fs_2048_tables.txt 2>&1
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py prove 256 > gpurun_out/r02n_trace_prove_256_tables.txt 2>&1
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify 2048 > gpurun_out/r02n_trace_verify_2048_tables.txt 2>&1
grep -v "^1 " gpurun_out/r02n_trace_verify_fs_2048_tables.txt
cat gpurun_out/r02n_trace_prove_256_tables.txt
grep -v "^1 " gpurun_out/r02n_trace_verify_2048_tables.txt
python bench.py --no-extras --steps 30 --warmup 5 > gpurun_out/r02n_bench_noextras.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/r02n_bench_noextras.json').read()); print(round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3))"
