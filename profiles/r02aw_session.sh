#!/bin/bash
# GPU session r02aw (clean full build of the final sources incl. bpp_points_precompute_direct): whole GPU suite, smoke, full bench line at N = 1,
# reference arm, launch list of the bench command.
set -x
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02aw_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02aw_pytest.txt
tail -4 gpurun_out/r02aw_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02aw_smoke.txt 2>&1; tail -1 gpurun_out/r02aw_smoke.txt
timeout 1800 python bench.py > gpurun_out/r02aw_bench_n1.json 2> gpurun_out/r02aw_bench_n1.err; echo "bench rc=$?"; tail -3 gpurun_out/r02aw_bench_n1.err
head -c 300 gpurun_out/r02aw_bench_n1.json
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02aw_bench_reference.json 2> gpurun_out/r02aw_bench_reference.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r02aw_launches_bench_steps2.csv python bench.py --steps 2 --warmup 3 --no-extras > gpurun_out/r02aw_ncu_bench.log 2>&1
python tools/launch_summary.py gpurun_out/r02aw_launches_bench_steps2.csv 140 > gpurun_out/r02aw_launch_summary.txt 2>&1; head -16 gpurun_out/r02aw_launch_summary.txt
