#!/bin/bash
# GPU session r02h: fused exponents in the direct kernel (A/B with BPP_NO_FUSE), launch bounds of verify_tables / wip_prove.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02h_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02h_pytest.txt
tail -3 gpurun_out/r02h_pytest.txt
for i in 1 2; do
  python bench.py --no-extras > gpurun_out/r02h_bench_fused_$i.json 2>/dev/null
  BPP_NO_FUSE=1 python bench.py --no-extras > gpurun_out/r02h_bench_nofuse_$i.json 2>/dev/null
done
for f in gpurun_out/r02h_bench_*.json; do python -c "
import json; d=json.loads(open('$f').read()); print('$f', round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3))"; done
BPP_TABLE_GB=80 timeout 600 python tools/run_case.py verify_fs 2048 5 > gpurun_out/r02h_verify_fs.txt 2>&1; tail -2 gpurun_out/r02h_verify_fs.txt
BPP_TABLE_GB=80 timeout 600 python tools/run_case.py prove 256 5 > gpurun_out/r02h_prove.txt 2>&1; tail -2 gpurun_out/r02h_prove.txt
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify_fs 2048 > gpurun_out/r02h_trace_verify_fs_2048_tables.txt 2>&1
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py prove 256 > gpurun_out/r02h_trace_prove_256_tables.txt 2>&1
grep -v "^1 " gpurun_out/r02h_trace_verify_fs_2048_tables.txt | awk '$4>25 || /^#/'
awk '$4>25 || /^#/' gpurun_out/r02h_trace_prove_256_tables.txt
