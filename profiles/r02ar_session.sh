#!/bin/bash
# r02ar: direct tables for ANY number of jobs (wide slicing + tree join); A/B with BPP_DIRECT_MIN_JOBS=32 (the old threshold)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_direct.py tests/test_gpu_straus.py tests/test_gpu_parity.py tests/test_gpu_session.py tests/test_gpu_rlc.py tests/test_gpu_hashed_pk.py -m gpu -x -q > gpurun_out/r02ar_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02ar_pytest.txt
tail -5 gpurun_out/r02ar_pytest.txt
for mj in 1 32; do
  export BPP_DIRECT_MIN_JOBS=$mj
  echo "== BPP_DIRECT_MIN_JOBS=$mj"
  BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py verify 1 2>&1 | grep -E "span"
  BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py verify_fs 1 2>&1 | grep -E "span"
  BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py prove 1 2>&1 | grep -E "span"
  BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py verify 16 2>&1 | grep -E "span"
  BPP_TABLE_GB=60 BPP_NM=64,64 timeout 600 python tools/trace_case.py verify 1 2>&1 | grep -E "span"
  BPP_TABLE_GB=60 BPP_NM=64,64 timeout 600 python tools/trace_case.py prove 1 2>&1 | grep -E "span"
done 2>&1 | tee gpurun_out/r02ar_ab.txt
unset BPP_DIRECT_MIN_JOBS
BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py verify 1 > gpurun_out/r02ar_trace_verify_64x16_single_tables.txt 2>&1; cat gpurun_out/r02ar_trace_verify_64x16_single_tables.txt
BPP_TABLE_GB=80 BPP_NM=64,16 timeout 600 python tools/trace_case.py prove 1 > gpurun_out/r02ar_trace_prove_64x16_single_tables.txt 2>&1; cat gpurun_out/r02ar_trace_prove_64x16_single_tables.txt
