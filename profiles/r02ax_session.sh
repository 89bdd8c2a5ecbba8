#!/bin/bash
# r02ax: C++ host mirror test incl. PublicKey::precompute
mkdir -p gpurun_out
python -m pytest tests/test_gpu_cpp_host.py -m gpu -x -q > gpurun_out/r02ax_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02ax_pytest.txt
tail -4 gpurun_out/r02ax_pytest.txt
bulletproofsplus_b200/bin/host_mirror_test 0 2>&1 | tail -8
