#!/bin/bash
# r02o: config 4 (8192 x (n=64, m=1)) with and without direct tables, c = 15 vs 16
mkdir -p gpurun_out
for gb in 0 6 12; do echo "== BPP_TABLE_GB=$gb"; BPP_NM=64,1 BPP_TABLE_GB=$gb timeout 600 python tools/run_case.py verify 8192 5 2>&1 | tail -3; done | tee gpurun_out/r02o_config4_tables.txt
for gb in 0 6; do echo "== 1024 proofs BPP_TABLE_GB=$gb"; BPP_NM=64,1 BPP_TABLE_GB=$gb timeout 600 python tools/run_case.py verify 1024 5 2>&1 | tail -3; done | tee -a gpurun_out/r02o_config4_tables.txt
BPP_NM=64,1 BPP_TABLE_GB=6 timeout 600 python tools/trace_case.py verify 8192 > gpurun_out/r02o_trace_config4.txt 2>&1; cat gpurun_out/r02o_trace_config4.txt
