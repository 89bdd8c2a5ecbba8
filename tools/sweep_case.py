"""The MSM sweep section of bench.py alone (development aid): device-resident ms per size for the three curves.
usage: sweep_case.py [max_log]"""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

args = argparse.Namespace(table_gb=0.0, batch=2048, msm_max_log=int(sys.argv[1]) if len(sys.argv) > 1 else 22)
bn = bench.Bench(args, torch, None, 0, 1, 0)
probe = bn.ctx.int_peak_probe()
out = bench.run_msm_sweep(bn, probe["imad32_per_s"])
for k, rows in out.items():
    print(k, json.dumps([(r["log2_n"], round(r["ms"], 3), r["path"]) for r in rows]))
