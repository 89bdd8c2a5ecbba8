#!/bin/bash
# r02an: thread-block cluster per proof in wip_prove_kernel for small batches (Ristretto objects only). A/B: BPP_NO_CLUSTER=1
mkdir -p gpurun_out
python -m pytest tests/test_gpu_parity.py tests/test_gpu_full_size.py tests/test_gpu_session.py tests/test_gpu_direct.py -m gpu -x -q > gpurun_out/r02an_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02an_pytest.txt
tail -5 gpurun_out/r02an_pytest.txt
for nc in 0 1; do
  if [ $nc = 1 ]; then export BPP_NO_CLUSTER=1; else unset BPP_NO_CLUSTER; fi
  echo "== BPP_NO_CLUSTER=$nc"
  BPP_NM=64,64 timeout 600 python tools/trace_case.py prove 1 2>&1 | grep -E "span|wip_prove"
  BPP_NM=64,16 timeout 600 python tools/trace_case.py prove 1 2>&1 | grep -E "span|wip_prove"
  BPP_NM=64,16 timeout 600 python tools/trace_case.py prove 16 2>&1 | grep -E "span|wip_prove"
  BPP_NM=64,1 timeout 600 python tools/trace_case.py prove 8 2>&1 | grep -E "span|wip_prove"
done 2>&1 | tee gpurun_out/r02an_ab.txt
