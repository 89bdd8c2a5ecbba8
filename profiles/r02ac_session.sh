#!/bin/bash
# r02ac: A/B of the in-library pipelining, device-resident (bench.py's per-proof-challenge section alone)
mkdir -p gpurun_out
for np in 0 1 0 1; do
  if [ $np = 1 ]; then export BPP_NO_PIPELINE=1; else unset BPP_NO_PIPELINE; fi
  echo "== BPP_NO_PIPELINE=$np"; timeout 600 python tools/fs_case.py 80 2>&1 | tail -1
done 2>&1 | tee gpurun_out/r02ac_ab.txt
