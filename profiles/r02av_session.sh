#!/bin/bash
# r02av: bpp_points_precompute_direct - parity tests and the MSM sweep with direct tables for reused point sets
mkdir -p gpurun_out
python -m pytest tests/test_gpu_direct.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02av_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02av_pytest.txt
tail -6 gpurun_out/r02av_pytest.txt
timeout 900 python tools/sweep_case.py 22 2>&1 | tail -3 | tee gpurun_out/r02av_sweep.txt
