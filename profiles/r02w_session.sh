#!/bin/bash
# r02w: Straus window sums for the per-call points (msm_straus.cuh) - Ristretto objects only; A/B with BPP_NO_STRAUS=1
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_gpu_session.py tests/test_gpu_full_size.py tests/test_gpu_two_contexts.py -m gpu -x -q > gpurun_out/r02w_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02w_pytest.txt
tail -3 gpurun_out/r02w_pytest.txt
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify 2048 > gpurun_out/r02w_trace_verify_2048_tables.txt 2>&1; cat gpurun_out/r02w_trace_verify_2048_tables.txt
BPP_NM=64,1 BPP_TABLE_GB=6 timeout 600 python tools/trace_case.py verify 8192 > gpurun_out/r02w_trace_config4.txt 2>&1; cat gpurun_out/r02w_trace_config4.txt
for ns in 0 1; do
  if [ $ns = 1 ]; then export BPP_NO_STRAUS=1; else unset BPP_NO_STRAUS; fi
  echo "== BPP_NO_STRAUS=$ns"
  BPP_NM=64,1 BPP_TABLE_GB=6 timeout 600 python tools/run_case.py verify 8192 5 2>&1 | tail -2 | head -1
  BPP_NM=64,1 BPP_TABLE_GB=6 timeout 600 python tools/run_case.py verify 1024 5 2>&1 | tail -2 | head -1
  BPP_NM=64,1 timeout 600 python tools/run_case.py verify 8192 5 2>&1 | tail -2 | head -1
  timeout 600 python tools/run_case.py verify 2048 5 2>&1 | tail -2 | head -1
  python bench.py --no-extras --steps 20 --warmup 3 > gpurun_out/r02w_bench_noextras_ns$ns.json 2>/dev/null
  python -c "
import json; d=json.loads(open('gpurun_out/r02w_bench_noextras_ns$ns.json').read()); print(round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3), round(d['e2e']['value']))"
done 2>&1 | tee gpurun_out/r02w_ab.txt
