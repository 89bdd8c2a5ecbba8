#!/bin/bash
# GPU session r02k: direct kernel with the shifted-scalar digit loop (1081 instructions per addition instead of 1342).
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02k_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02k_pytest.txt
tail -3 gpurun_out/r02k_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02k_smoke.txt 2>&1; tail -1 gpurun_out/r02k_smoke.txt
for i in 1 2; do python bench.py --no-extras > gpurun_out/r02k_bench_noextras_$i.json 2>/dev/null; done
python bench.py --no-extras --table-gb 48 > gpurun_out/r02k_bench_noextras_48gb.json 2>/dev/null
for f in gpurun_out/r02k_bench_*.json; do python -c "
import json; d=json.loads(open('$f').read()); print('$f', round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3), round(d['e2e']['value']))"; done
BPP_TABLE_GB=80 timeout 600 python tools/run_case.py verify_fs 2048 5 > gpurun_out/r02k_verify_fs.txt 2>&1; tail -2 gpurun_out/r02k_verify_fs.txt
BPP_TABLE_GB=80 timeout 600 python tools/run_case.py prove 256 5 > gpurun_out/r02k_prove.txt 2>&1; tail -2 gpurun_out/r02k_prove.txt
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify_fs 2048 > gpurun_out/r02k_trace_verify_fs_2048_tables.txt 2>&1
grep -v "^1 " gpurun_out/r02k_trace_verify_fs_2048_tables.txt | awk '$4>25 || /^#/'
