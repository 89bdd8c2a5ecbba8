#!/bin/bash
# r02ba: whole GPU suite + smoke on the final commit
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02ba_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02ba_pytest.txt
tail -4 gpurun_out/r02ba_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02ba_smoke.txt 2>&1; tail -1 gpurun_out/r02ba_smoke.txt
