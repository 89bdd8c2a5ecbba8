#!/bin/bash
# r02l: direct kernel with ONE inlined addition in phase 1 (instruction-cache fix of the r02k regression)
mkdir -p gpurun_out
python -m pytest tests/test_gpu_direct.py tests/test_gpu_parity.py tests/test_gpu_session.py -m gpu -x -q > gpurun_out/r02l_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02l_pytest.txt
python bench.py --no-extras --steps 30 --warmup 5 > gpurun_out/r02l_bench_noextras_1.json 2> gpurun_out/r02l_bench_1.err
python bench.py --no-extras --steps 30 --warmup 5 > gpurun_out/r02l_bench_noextras_2.json 2> gpurun_out/r02l_bench_2.err
tail -3 gpurun_out/r02l_pytest.txt
python - <<'PY'
import json
for f in ("1","2"):
    try:
        d=json.loads(open(f"gpurun_out/r02l_bench_noextras_{f}.json").read().strip().splitlines()[-1])
        print(f, d["value"], d["ms_per_step"], d["roofline"].get("kernel_ms"), d["e2e"]["value"])
    except Exception as e: print(f, "ERR", e)
PY
