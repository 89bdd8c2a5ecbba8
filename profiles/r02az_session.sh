#!/bin/bash
# r02az: the per-proof-challenge bench section after Context.close() was added for the second context
mkdir -p gpurun_out
timeout 600 python tools/fs_case.py 80 2>&1 | tail -2 | tee gpurun_out/r02az_fs.txt
python -m pytest tests/test_gpu_two_contexts.py -m gpu -x -q 2>&1 | tail -2
