#!/bin/bash
# r02au: switches test with BPP_DIRECT_MIN_JOBS and the few-jobs calls in its workload
mkdir -p gpurun_out
python -m pytest tests/test_gpu_switches.py -m gpu -x -q > gpurun_out/r02au_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02au_pytest.txt
tail -6 gpurun_out/r02au_pytest.txt
