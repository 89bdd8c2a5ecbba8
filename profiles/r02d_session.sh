#!/bin/bash
# GPU session r02d: A/B of the Karatsuba field multiply (libbpp_b200_kara.so, BPP_LIB) against the 64-product multiply:
# isolated multiply rate (fe_probe), parity subset, headline step twice each in alternation; then one ncu --set full capture of
# the per-proof-challenge verify (tables, sort, accumulate kernels) with the default library.
set -x
mkdir -p gpurun_out
K=$PWD/bulletproofsplus_b200/libbpp_b200_kara.so
python tools/fe_probe.py > gpurun_out/r02d_fe_probe_base.txt 2>&1
BPP_LIB=$K python tools/fe_probe.py > gpurun_out/r02d_fe_probe_kara.txt 2>&1
BPP_LIB=$K timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02d_pytest_kara.txt 2>&1; echo "rc=$?" >> gpurun_out/r02d_pytest_kara.txt
tail -3 gpurun_out/r02d_pytest_kara.txt
for i in 1 2; do
  python bench.py --no-extras > gpurun_out/r02d_bench_base_$i.json 2>/dev/null
  BPP_LIB=$K python bench.py --no-extras > gpurun_out/r02d_bench_kara_$i.json 2>/dev/null
done
for f in gpurun_out/r02d_bench_*.json; do python -c "
import json,sys; d=json.loads(open('$f').read()); print('$f', round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3))"; done
BPP_LIB=$K python tools/run_case.py msm 22 3 > gpurun_out/r02d_msm22_kara.txt 2>&1; python tools/run_case.py msm 22 3 > gpurun_out/r02d_msm22_base.txt 2>&1
BPP_LIB=$K python tools/run_case.py prove 256 3 > gpurun_out/r02d_prove_kara.txt 2>&1; python tools/run_case.py prove 256 3 > gpurun_out/r02d_prove_base.txt 2>&1
tail -2 gpurun_out/r02d_msm22_*.txt gpurun_out/r02d_prove_*.txt
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'verify_tables_kernel|msm_sort_job_kernel|verify_inverses_kernel|msm_reduce_chunks_kernel|msm_accumulate_kernel' --launch-skip 24 --launch-count 8 -o gpurun_out/r02d_verify_fs python tools/run_case.py verify_fs 2048 1 > gpurun_out/r02d_ncu.log 2>&1
ncu -i gpurun_out/r02d_verify_fs.ncu-rep --page details > gpurun_out/r02d_verify_fs_kernels_ncu_full_details.txt 2>&1
grep -E "^  [a-z_]+.*\(|Duration|Registers Per|Issue Slots Busy|SM Busy|Achieved Occupancy|Executed Ipc Active" gpurun_out/r02d_verify_fs_kernels_ncu_full_details.txt | head -60
