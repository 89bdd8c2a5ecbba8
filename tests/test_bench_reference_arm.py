"""`bench.py --impl reference` (the CPU arm the driver runs beside the GPU arm) needs no GPU: one short run here checks the
contract of its JSON line — the metric and config of the GPU arm, `impl`, a `cpu_baseline` describing the run, and an `e2e`
that repeats the value with zero bytes copied."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "0", "--ref-proofs", "1"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.splitlines() if l.startswith("{")]
    assert len(lines) == 1, r.stdout[-2000:]
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "range_proof_verifies_per_sec_n64_m16" and d["unit"] == "proofs/s"
    assert d["higher_is_better"] is True and d["n_gpus"] == 1 and d["steps"] == 1
    assert d["value"] > 0 and abs(d["value"] - d["e2e"]["value"]) < 1e-9
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    cb = d["cpu_baseline"]
    assert cb["kind"] == "port" and cb["cores"] == 1 and cb["value"] == d["value"] and "sample" in cb
    assert cb["b2_fair_cpu"]["value"] > cb["value"]  # the bucket-MSM verifier beats the literal loop
    assert "n=64" in d["config"]["workload"]
