#!/bin/bash
# GPU session r02e: direct fixed-base tables (bpp_pk_precompute), hashed generators, batched window-table build, bpp_pk_msm.
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02e_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02e_pytest.txt
tail -6 gpurun_out/r02e_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02e_smoke.txt 2>&1; tail -1 gpurun_out/r02e_smoke.txt
for gb in 0 12 48 80; do
  BPP_TABLE_GB=$gb timeout 600 python tools/run_case.py verify 2048 5 > gpurun_out/r02e_verify_tables_${gb}gb.txt 2>&1; tail -3 gpurun_out/r02e_verify_tables_${gb}gb.txt
done
BPP_TABLE_GB=48 timeout 600 python tools/run_case.py prove 256 5 > gpurun_out/r02e_prove_tables_48gb.txt 2>&1; tail -2 gpurun_out/r02e_prove_tables_48gb.txt
BPP_TABLE_GB=48 timeout 600 python tools/trace_case.py verify 2048 > gpurun_out/r02e_trace_verify_2048_tables.txt 2>&1
BPP_TABLE_GB=48 timeout 600 python tools/trace_case.py verify_fs 2048 > gpurun_out/r02e_trace_verify_fs_2048_tables.txt 2>&1
BPP_TABLE_GB=48 timeout 600 python tools/trace_case.py prove 256 > gpurun_out/r02e_trace_prove_256_tables.txt 2>&1
timeout 1800 python bench.py > gpurun_out/r02e_bench_n1.json 2> gpurun_out/r02e_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02e_bench_n1.err
head -c 400 gpurun_out/r02e_bench_n1.json
BPP_TABLE_GB=48 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'direct_accumulate_kernel' --launch-skip 3 --launch-count 1 -o gpurun_out/r02e_direct python tools/run_case.py verify 2048 1 > gpurun_out/r02e_ncu.log 2>&1
ncu -i gpurun_out/r02e_direct.ncu-rep --page details > gpurun_out/r02e_direct_accumulate_ncu_full_details.txt 2>&1
ncu -i gpurun_out/r02e_direct.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[0]
for name in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','sm__inst_executed_pipe_fmaheavy.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__t_sector_hit_rate.pct'):
    if name in h: print(name, rows[1][h.index(name)], rows[2][h.index(name)] if len(rows)>2 else '')
" > gpurun_out/r02e_direct_raw_metrics.txt 2>&1; cat gpurun_out/r02e_direct_raw_metrics.txt
