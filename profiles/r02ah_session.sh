#!/bin/bash
# r02ah: MSM sweep with the Weierstrass field multiply as a function (compare profiles/r02x_bench_n1.json)
mkdir -p gpurun_out
timeout 900 python tools/sweep_case.py 22 2>&1 | tail -3 | tee gpurun_out/r02ah_sweep.txt
