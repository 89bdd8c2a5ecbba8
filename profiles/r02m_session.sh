#!/bin/bash
# r02m (1 GPU): secp256k1 special-form field; full GPU suite, smoke, full bench line, reference arm, 146 GB tables (c = 16) trial
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r02m_pytest.txt 2>&1; echo "pytest rc=$?" >> gpurun_out/r02m_pytest.txt
tail -3 gpurun_out/r02m_pytest.txt
timeout 300 python __graft_entry__.py smoke > gpurun_out/r02m_smoke.txt 2>&1; tail -1 gpurun_out/r02m_smoke.txt
timeout 1500 python bench.py > gpurun_out/r02m_bench_n1.json 2> gpurun_out/r02m_bench_n1.err; echo "bench rc=$?"
timeout 600 python bench.py --impl reference > gpurun_out/r02m_bench_reference.json 2> gpurun_out/r02m_bench_reference.err; echo "ref rc=$?"
timeout 600 python bench.py --no-extras --table-gb 150 > gpurun_out/r02m_bench_noextras_c16.json 2> gpurun_out/r02m_bench_c16.err; echo "c16 rc=$?"
BPP_TABLE_GB=80 timeout 600 python tools/trace_case.py verify_fs 2048 > gpurun_out/r02m_trace_verify_fs_2048_tables.txt 2>&1
python - <<'PY'
import json
for f in ("r02m_bench_n1","r02m_bench_noextras_c16"):
    try:
        d=json.loads(open(f"gpurun_out/{f}.json").read().strip().splitlines()[-1])
        print(f, round(d["value"]), round(d["ms_per_step"],3), round(d["roofline"]["avg_launch_ms"],3), round(d["e2e"]["value"]), d["config"].get("fixed_base_tables"), d.get("extras_errors"))
    except Exception as e: print(f, "ERR", e)
PY
