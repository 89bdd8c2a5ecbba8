mkdir -p gpurun_out
python -m pytest tests/test_gpu_direct.py tests/test_gpu_switches.py -m gpu -x -q > gpurun_out/r02aj_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02aj_pytest.txt
tail -12 gpurun_out/r02aj_pytest.txt
