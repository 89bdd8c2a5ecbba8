#!/bin/bash
# r02ak: timelines of config 3's single-proof calls (n = 64, m = 64): where the 0.87 ms verify / 2.83 ms prove go
mkdir -p gpurun_out
BPP_NM=64,64 timeout 600 python tools/trace_case.py verify 1 > gpurun_out/r02ak_trace_verify_64x64_single.txt 2>&1; cat gpurun_out/r02ak_trace_verify_64x64_single.txt
BPP_NM=64,64 timeout 600 python tools/trace_case.py prove 1 > gpurun_out/r02ak_trace_prove_64x64_single.txt 2>&1; cat gpurun_out/r02ak_trace_prove_64x64_single.txt
