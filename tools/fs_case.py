"""Device-resident timing of the per-proof-challenge verification section of bench.py alone (development aid):
usage: fs_case.py [table_gb]   env: BPP_NO_STRAUS=1 for A/B."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import bench

gb = float(sys.argv[1]) if len(sys.argv) > 1 else 80.0
args = argparse.Namespace(table_gb=gb, batch=2048)
bn = bench.Bench(args, torch, None, 0, 1, 0)
pk = bn.api.PublicKey.new(bench.N_BITS * bench.M_AGG, 0, bn.ctx)
if gb > 0:
    print("direct tables c =", pk.precompute(int(gb * 2**30)), flush=True)
out = bench.run_fs_verify(bn, pk, args.batch)
fs = out["verify_per_proof_challenges"]
print(json.dumps({"value": fs["value"], "ms": fs["ms_per_step"], "e2e": fs["e2e"]["value"], "two_contexts": fs.get("two_contexts", {}).get("value"),
                  "rlc": out["verifies_per_sec_rlc_batch"]["value"]}))
