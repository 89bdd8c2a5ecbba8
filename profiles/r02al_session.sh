#!/bin/bash
# r02al: latency tiers up to 262144 buckets (was 65536): single-proof calls at n = 64, m = 64, WIP 2^12, MSM sweep
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_curves.py -m gpu -x -q 2>&1 | tail -2
BPP_NM=64,64 timeout 600 python tools/trace_case.py verify 1 2>&1 | grep -E "span|k32|k4|msm_accumulate" 
BPP_NM=64,64 timeout 600 python tools/trace_case.py prove 1 2>&1 | grep -E "span|k32|k4|msm_accumulate"
timeout 900 python tools/sweep_case.py 22 2>&1 | tail -3 | tee gpurun_out/r02al_sweep.txt
