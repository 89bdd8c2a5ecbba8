mkdir -p gpurun_out
python -m pytest tests/test_gpu_straus.py tests/test_gpu_two_contexts.py tests/test_gpu_curves.py -m gpu -x -q > gpurun_out/r02aa_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02aa_pytest.txt
tail -30 gpurun_out/r02aa_pytest.txt
