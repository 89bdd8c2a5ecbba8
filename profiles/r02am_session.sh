#!/bin/bash
# r02am: last check of the library as it sits in the tree (clean build after the final source change, LAT_MAX_KEYS):
# whole GPU suite, smoke, headline line without extras.
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r02am_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02am_pytest.txt
tail -4 gpurun_out/r02am_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02am_smoke.txt 2>&1; tail -1 gpurun_out/r02am_smoke.txt
python bench.py --no-extras > gpurun_out/r02am_bench_noextras.json 2>/dev/null
python -c "
import json; d=json.loads(open('gpurun_out/r02am_bench_noextras.json').read()); print(round(d['value']), round(d['ms_per_step'],3), round(d['roofline']['avg_launch_ms'],3), round(d['e2e']['value']), d['gpu_launches'])"
