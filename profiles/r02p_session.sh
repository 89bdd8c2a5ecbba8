#!/bin/bash
# r02p: Ed25519 full addition as a real function (reduction kernels) - A/B against r02n/r02o traces
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02p_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02p_pytest.txt
tail -2 gpurun_out/r02p_pytest.txt
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify 2048 > gpurun_out/r02p_trace_verify_2048_tables.txt 2>&1
BPP_NM=64,1 BPP_TABLE_GB=6 timeout 600 python tools/trace_case.py verify 8192 > gpurun_out/r02p_trace_config4.txt 2>&1
timeout 600 python tools/trace_case.py msm 22 > gpurun_out/r02p_trace_msm_22.txt 2>&1
timeout 600 python tools/trace_case.py msm 10 > gpurun_out/r02p_trace_msm_10.txt 2>&1
cat gpurun_out/r02p_trace_verify_2048_tables.txt gpurun_out/r02p_trace_config4.txt gpurun_out/r02p_trace_msm_22.txt gpurun_out/r02p_trace_msm_10.txt
for gb in 6; do BPP_NM=64,1 BPP_TABLE_GB=$gb timeout 600 python tools/run_case.py verify 8192 5 2>&1 | tail -2; done
python bench.py --no-extras --steps 30 --warmup 5 > gpurun_out/r02p_bench_noextras.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/r02p_bench_noextras.json').read()); print(round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3))"
