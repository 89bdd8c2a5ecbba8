#!/bin/bash
# GPU session r02x (clean full build incl. the Straus path for per-proof points): whole GPU suite, smoke, full bench line at N = 1,
# reference arm, launch list of the bench command.
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02x_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02x_pytest.txt
tail -4 gpurun_out/r02x_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02x_smoke.txt 2>&1; tail -1 gpurun_out/r02x_smoke.txt
timeout 1800 python bench.py > gpurun_out/r02x_bench_n1.json 2> gpurun_out/r02x_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02x_bench_n1.err
head -c 300 gpurun_out/r02x_bench_n1.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02x_bench_reference.json 2> gpurun_out/r02x_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02x_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02x_ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r02x_launches_bench_steps2.csv 140 > gpurun_out/r02x_launch_summary.txt 2>&1; head -16 gpurun_out/r02x_launch_summary.txt
BPP_TABLE_GB=80 timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'direct_accumulate_kernel|msm_straus_kernel|msm_combine4_kernel' --launch-skip 6 --launch-count 3 -o gpurun_out/r02x_verify_kernels python tools/run_case.py verify 2048 1 > gpurun_out/r02x_ncu.log 2>&1
ncu -i gpurun_out/r02x_verify_kernels.ncu-rep --page details > gpurun_out/r02x_verify_kernels_ncu_full_details.txt 2>&1
ncu -i gpurun_out/r02x_verify_kernels.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
names=('gpu__time_duration.sum','launch__registers_per_thread','sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','dram__bytes_write.sum','lts__t_sector_hit_rate.pct','smsp__thread_inst_executed_per_inst_executed.ratio','sm__warps_active.avg.pct_of_peak_sustained_active')
for r in rows[2:]:
    print(r[h.index('Kernel Name')][:70])
    for n in names:
        if n in h: print('   ', n, r[h.index(n)], rows[1][h.index(n)])
" > gpurun_out/r02x_verify_kernels_raw_metrics.txt 2>&1; cat gpurun_out/r02x_verify_kernels_raw_metrics.txt
