#!/bin/bash
# r02aq: secondary-curve parity incl. the cluster prover shapes; switches test incl. BPP_NO_CLUSTER
mkdir -p gpurun_out
python -m pytest tests/test_gpu_curves.py tests/test_gpu_switches.py -m gpu -x -q > gpurun_out/r02aq_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02aq_pytest.txt
tail -6 gpurun_out/r02aq_pytest.txt
