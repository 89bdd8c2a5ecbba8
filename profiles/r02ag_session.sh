#!/bin/bash
# r02ag: base-field multiply of the Weierstrass curves as a function - parity tests and MSM timings (compare r02x sweep)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_curves.py tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/r02ag_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02ag_pytest.txt
tail -3 gpurun_out/r02ag_pytest.txt
for cv in 1 2; do for lg in 10 14 16 20 22; do timeout 600 python tools/run_case.py msm $lg 3 $cv 2>&1 | tail -2 | head -1; done; done | tee gpurun_out/r02ag_msm.txt
