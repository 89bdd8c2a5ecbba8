#!/bin/bash
# GPU session r02c: four-warp Horner combine, window-grouped digit scatter, per-proof tables restructure, commit_u64, fresh nonces.
# Tests -> timelines -> full bench line -> reference arm -> ncu launch list + one --set full capture of the bench's top kernels.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02c_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02c_pytest.txt
tail -5 gpurun_out/r02c_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02c_smoke.txt 2>&1; tail -2 gpurun_out/r02c_smoke.txt
for c in "verify 2048" "verify_fs 2048" "prove 256"; do timeout 300 python tools/trace_case.py $c > gpurun_out/r02c_trace_${c// /_}.txt 2>&1; done
for lg in 10 14 18 20 22; do timeout 300 python tools/trace_case.py msm $lg > gpurun_out/r02c_trace_msm_$lg.txt 2>&1; done
timeout 300 python tools/trace_case.py msm 10 tables > gpurun_out/r02c_trace_msm_10_tables.txt 2>&1
timeout 1500 python bench.py > gpurun_out/r02c_bench_n1.json 2> gpurun_out/r02c_bench_n1.err; echo "bench rc=$?"
tail -3 gpurun_out/r02c_bench_n1.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02c_bench_reference.json 2> gpurun_out/r02c_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file gpurun_out/r02c_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02c_ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r02c_launches_bench_steps2.csv | head -30
