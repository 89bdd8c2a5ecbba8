#!/bin/bash
# GPU session r02q (clean full build of the final sources): whole GPU suite, smoke, full bench line at N = 1, reference arm,
# launch list, ncu capture of the final direct kernel and of the per-proof challenge-table kernel (its stall reasons).
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02q_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02q_pytest.txt
tail -4 gpurun_out/r02q_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02q_smoke.txt 2>&1; tail -1 gpurun_out/r02q_smoke.txt
timeout 1800 python bench.py > gpurun_out/r02q_bench_n1.json 2> gpurun_out/r02q_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02q_bench_n1.err
head -c 300 gpurun_out/r02q_bench_n1.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02q_bench_reference.json 2> gpurun_out/r02q_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02q_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02q_ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r02q_launches_bench_steps2.csv 140 | head -24
BPP_TABLE_GB=80 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'direct_accumulate_kernel' --launch-skip 3 --launch-count 1 -o gpurun_out/r02q_direct python tools/run_case.py verify 2048 1 > gpurun_out/r02q_ncu.log 2>&1
ncu -i gpurun_out/r02q_direct.ncu-rep --page details > gpurun_out/r02q_direct_accumulate_ncu_full_details.txt 2>&1
ncu -i gpurun_out/r02q_direct.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[0]
for name in ('dram__bytes_read.sum','dram__bytes_write.sum','gpu__time_duration.sum','smsp__issue_active.avg.pct_of_peak_sustained_active','launch__registers_per_thread','dram__throughput.avg.pct_of_peak_sustained_elapsed','lts__t_sector_hit_rate.pct','sm__throughput.avg.pct_of_peak_sustained_elapsed','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio'):
    if name in h: print(name, rows[1][h.index(name)], rows[2][h.index(name)] if len(rows)>2 else '')
" > gpurun_out/r02q_direct_raw_metrics.txt 2>&1; cat gpurun_out/r02q_direct_raw_metrics.txt

BPP_TABLE_GB=80 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'verify_tables_kernel|verify_inverses_kernel|wip_prove_kernel' --launch-skip 2 --launch-count 2 -o gpurun_out/r02q_scalar_kernels python tools/run_case.py verify_fs 2048 1 > gpurun_out/r02q_ncu2.log 2>&1
ncu -i gpurun_out/r02q_scalar_kernels.ncu-rep --page details > gpurun_out/r02q_scalar_kernels_ncu_full_details.txt 2>&1
ncu -i gpurun_out/r02q_scalar_kernels.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin))
h=rows[0]
want=[i for i,n in enumerate(h) if n in ('Kernel Name','gpu__time_duration.sum','launch__registers_per_thread','smsp__issue_active.avg.pct_of_peak_sustained_active') or 'warp_issue_stalled' in n and n.endswith('_per_warp_active.pct')]
for r in [h]+rows[2:]:
    print(' | '.join(r[i][:60] for i in want))
" > gpurun_out/r02q_scalar_kernels_stalls.txt 2>&1; head -c 3000 gpurun_out/r02q_scalar_kernels_stalls.txt
