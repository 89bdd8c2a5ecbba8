#!/bin/bash
# r02ab: in-library two-half pipelining of per-proof-challenge verification (Ristretto objects + bpp.o only)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_straus.py tests/test_gpu_session.py tests/test_gpu_two_contexts.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r02ab_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02ab_pytest.txt
tail -15 gpurun_out/r02ab_pytest.txt
for np in 0 1; do
  if [ $np = 1 ]; then export BPP_NO_PIPELINE=1; else unset BPP_NO_PIPELINE; fi
  echo "== BPP_NO_PIPELINE=$np"
  BPP_TABLE_GB=80 timeout 600 python tools/run_case.py verify_fs 2048 8 2>&1 | tail -2 | head -1
  BPP_TABLE_GB=80 timeout 600 python tools/run_case.py verify_fs 4096 4 2>&1 | tail -2 | head -1
done 2>&1 | tee gpurun_out/r02ab_ab.txt
