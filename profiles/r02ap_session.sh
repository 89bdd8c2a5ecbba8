#!/bin/bash
# r02ap: timelines of single-proof verification with per-proof challenges (n = 64, m = 16 and m = 64)
mkdir -p gpurun_out
BPP_NM=64,16 timeout 600 python tools/trace_case.py verify_fs 1 > gpurun_out/r02ap_trace_verify_fs_64x16_single.txt 2>&1; cat gpurun_out/r02ap_trace_verify_fs_64x16_single.txt
BPP_NM=64,64 timeout 600 python tools/trace_case.py verify_fs 1 > gpurun_out/r02ap_trace_verify_fs_64x64_single.txt 2>&1; grep -E "span|verify_|decode|combine|k32|k4" gpurun_out/r02ap_trace_verify_fs_64x64_single.txt
