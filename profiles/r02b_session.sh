#!/bin/bash
# GPU session r02b: parity after fe_sqr / combine / level-reduce / sort routing (+ the no-alloc entry point); timelines
# (bpp_trace) of the verify / prove steps and of single MSMs over the size sweep; a bench line without extras.
set -x
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/r02b_pytest_noalloc_alone.txt 2>&1; echo "rc=$?" >> gpurun_out/r02b_pytest_noalloc_alone.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02b_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02b_pytest.txt
tail -5 gpurun_out/r02b_pytest.txt gpurun_out/r02b_pytest_noalloc_alone.txt
for c in "verify 2048" "verify_fs 2048" "prove 256"; do timeout 300 python tools/trace_case.py $c > gpurun_out/r02b_trace_${c// /_}.txt 2>&1; done
for lg in 10 12 14 16 18 20 22; do
  timeout 300 python tools/trace_case.py msm $lg > gpurun_out/r02b_trace_msm_$lg.txt 2>&1
  [ $lg -le 16 ] && timeout 300 python tools/trace_case.py msm $lg tables > gpurun_out/r02b_trace_msm_${lg}_tables.txt 2>&1
done
timeout 600 python bench.py --no-extras > gpurun_out/r02b_bench_noextras.json 2> gpurun_out/r02b_bench_noextras.err
head -c 600 gpurun_out/r02b_bench_noextras.json
