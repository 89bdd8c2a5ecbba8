#!/bin/bash
# r02s: latency setting of the bucket tiers for small MSMs (<= 16384 buckets): Ristretto objects only
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_direct.py tests/test_gpu_full_size.py -m gpu -x -q > gpurun_out/r02s_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02s_pytest.txt
tail -2 gpurun_out/r02s_pytest.txt
for lg in 10 12 13 14 16; do timeout 300 python tools/trace_case.py msm $lg tables > gpurun_out/r02s_trace_msm_${lg}_tables.txt 2>&1; head -2 gpurun_out/r02s_trace_msm_${lg}_tables.txt | tail -1; grep -E "k32|k4|msm_accumulate_kernel" gpurun_out/r02s_trace_msm_${lg}_tables.txt; done
for lg in 10 12 14; do timeout 300 python tools/trace_case.py msm $lg generic > gpurun_out/r02s_trace_msm_${lg}_generic.txt 2>&1; head -2 gpurun_out/r02s_trace_msm_${lg}_generic.txt | tail -1; grep -E "k32|k4|msm_accumulate_kernel" gpurun_out/r02s_trace_msm_${lg}_generic.txt; done
python bench.py --no-extras --steps 20 --warmup 3 > gpurun_out/r02s_bench_noextras.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/r02s_bench_noextras.json').read()); print(round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3))"
