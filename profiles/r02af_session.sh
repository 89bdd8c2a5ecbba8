#!/bin/bash
# r02af: ncu --set full of msm_accumulate_kernel on the two secondary curves (2^20-point MSM): registers, spills, pipe use
mkdir -p gpurun_out
for cv in 1 2; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:'msm_accumulate_kernel' --launch-skip 2 --launch-count 1 -o gpurun_out/r02af_acc_curve$cv python tools/run_case.py msm 20 3 $cv > gpurun_out/r02af_ncu_$cv.log 2>&1
  ncu -i gpurun_out/r02af_acc_curve$cv.ncu-rep --page details > gpurun_out/r02af_msm_accumulate_curve${cv}_ncu_full_details.txt 2>&1
  ncu -i gpurun_out/r02af_acc_curve$cv.ncu-rep --page raw --csv 2>/dev/null | python -c "
import csv,sys
rows=list(csv.reader(sys.stdin)); h=rows[0]
names=('gpu__time_duration.sum','launch__registers_per_thread','sm__throughput.avg.pct_of_peak_sustained_elapsed','smsp__issue_active.avg.pct_of_peak_sustained_active','dram__bytes_read.sum','lts__t_sector_hit_rate.pct','smsp__thread_inst_executed_per_inst_executed.ratio','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum','launch__occupancy_limit_registers','smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio','smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio','smsp__average_warps_issue_stalled_wait_per_issue_active.ratio','smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio','sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active','sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active','local_load_spilling','derived__memory_l1_wavefronts_shared_excessive')
for r in rows[2:]:
    print(r[h.index('Kernel Name')][:90])
    for n in names:
        if n in h: print('   ', n, r[h.index(n)], rows[1][h.index(n)])
" | tee gpurun_out/r02af_msm_accumulate_curve${cv}_raw_metrics.txt
done
